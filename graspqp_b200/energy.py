"""calculate_energy — graspqp/src/graspqp/core/energy.py:6-66 with the reference's signature and dict order.
Every heavy term is a kernel behind an autograd bridge (ops.py); the few element-wise lines are the reference's own."""
import torch


def calculate_energy(hand_model, object_model, energy_fnc=None, energy_names=(), method="gendexgrasp", svd_gain=0.1):
    losses = {}
    distance, contact_normal = object_model.cal_distance(hand_model.contact_points)
    if method == "dexgraspnet":
        E_dis = torch.sum(distance.abs(), dim=-1, dtype=torch.float)
    elif method == "gendexgrasp":  # energy.py:24-28
        E_dis = ((1 - torch.sum((-contact_normal) * hand_model.contact_normals, dim=-1)).exp() * distance.abs()).sum(-1)
    else:
        raise ValueError(f"Unknown method: {method}")
    losses["E_dis"] = E_dis
    E_fc, _lambda = energy_fnc(contact_pts=hand_model.contact_points, contact_normals=contact_normal, sdf=distance,
                               cog=object_model.cog, with_solution=True, svd_gain=svd_gain)
    losses["E_fc"] = E_fc
    th = hand_model.hand_pose[:, 9:]
    losses["E_joints"] = torch.sum((th > hand_model.joints_upper) * (th - hand_model.joints_upper), dim=-1) + torch.sum(
        (th < hand_model.joints_lower) * (hand_model.joints_lower - th), dim=-1)
    d = hand_model.cal_distance(object_model.surface_points_tensor * object_model.object_scale)
    losses["E_pen"] = torch.relu(d).sum(-1)  # energy.py:60-61
    losses["E_spen"] = hand_model.self_penetration()
    if "E_prior" in energy_names:  # energy.py:68-74: the grasp axis should point down
        forward_axis = (hand_model.global_rotation @ hand_model.grasp_axis.view(1, -1, 1)).view(-1, 3)
        axis_prior = torch.tensor([0, 0, -1], dtype=torch.float, device=forward_axis.device).view(1, 3)
        losses["E_prior"] = 1 - torch.sum(forward_axis * axis_prior, dim=-1)
    if "E_wall" in energy_names:   # energy.py:75-77: nothing below the table plane z = 0
        losses["E_wall"] = hand_model.get_surface_points()[..., -1].clamp(max=0.0).abs().sum(-1)
    if "E_manipulativity" in energy_names:
        raise NotImplementedError("E_manipulativity: scripts/fit.py's weight_dict has no entry for it (the term raises there)")
    return losses
