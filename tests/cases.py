"""Parity cases shared by the GPU tests (libgraspqp_b200.so + torch CUDA tensors).  Each takes (lib, be)."""
import ctypes

import numpy as np
import torch

from graspqp_b200 import _lib, engine, hand_pack, synth
from oracle import qp as oqp
from oracle import sdf as osdf
from oracle.hand import HandOracle
from oracle.pipeline import FitLoopOracle
import tol


def sdf_forward(lib, be, pts, fv):
    P = len(pts)
    dp, df = be.from_host(pts, np.float32), be.from_host(fv.reshape(-1, 9), np.float32)
    d2, sg = be.empty((P,), np.float32), be.empty((P,), np.int32)
    nr, cl = be.empty((P, 3), np.float32), be.empty((P, 3), np.float32)
    _lib.check(lib.gqp_sdf_forward(be.ptr(dp), be.ptr(df), P, len(fv), be.ptr(d2), be.ptr(sg), be.ptr(nr), be.ptr(cl),
                                   be.stream), lib)
    return [be.to_host(a) for a in (d2, sg, nr, cl)]


def mesh_sdf_forward(lib, be, pts, fv):
    P = len(pts)
    fvh = np.ascontiguousarray(fv.reshape(-1, 9), np.float32)
    m = ctypes.c_void_p()
    _lib.check(lib.gqp_mesh_create(fvh.ctypes.data, len(fvh), ctypes.byref(m)), lib)
    dp = be.from_host(pts, np.float32)
    d2, sg, fi = be.empty((P,), np.float32), be.empty((P,), np.int32), be.empty((P,), np.int32)
    nr, cl = be.empty((P, 3), np.float32), be.empty((P, 3), np.float32)
    _lib.check(lib.gqp_mesh_sdf_forward(m, be.ptr(dp), P, be.ptr(d2), be.ptr(sg), be.ptr(nr), be.ptr(cl), be.ptr(fi),
                                        be.stream), lib)
    out = [be.to_host(a) for a in (d2, sg, nr, cl, fi)]
    lib.gqp_mesh_destroy(m)
    return out


def query_points(fv, P, seed=0, spread=0.15):
    rng = np.random.default_rng(seed)
    pts = rng.uniform(-spread, spread, (P, 3)).astype(np.float32)
    k = P // 4
    pts[:k] = (fv.reshape(-1, 3)[rng.integers(0, len(fv) * 3, k)] * (1 + rng.normal(size=(k, 1)) * 0.02)).astype(np.float32)
    return pts


def check_sdf_bit_exact(lib, be, kind, nf, P, seed=0):
    v, f = synth.make_mesh(kind, nf, seed + 1)
    fv = synth.face_vertices(v, f)
    pts = query_points(fv, P, seed)
    want = osdf.compute_sdf_port(pts, fv, with_face=True)
    got = sdf_forward(lib, be, pts, fv)
    for g, w in zip(got, want[:4]):
        np.testing.assert_array_equal(g, w)
    got = mesh_sdf_forward(lib, be, pts, fv)
    for g, w in zip(got, want):
        np.testing.assert_array_equal(g, w)


def fk_state(lib, be, hand, pack, pose):
    B, L = len(pose), len(pack["mesh_frame"])
    dpose = be.from_host(pose, np.float32)
    tf = be.empty((B, L, 12), np.float32)
    _lib.check(lib.gqp_fk_forward(hand, be.ptr(dpose), B, be.ptr(tf), be.stream), lib)
    return dpose, tf


def check_hand_distance(lib, be, B=64, n_pts=2500, seed=5, hand="allegro"):
    """culled E_pen / gradient == un-culled at the reference's sample count; distances vs the oracle on a subset."""
    pack = hand_pack.load_pack(hand)
    allegro = hand == "allegro"
    hand = hand_pack.create_hand(lib, pack)
    v, f = synth.make_mesh("hull", 2000, seed)
    sp = synth.sample_surface(v, f, n_pts, seed=seed, oversample=2)
    pose, idx = synth.random_grasp_init(pack, [v], B, 12, seed)
    # push the hand into the object: grippers have a wide open palm, they need to go deeper to touch anything
    pose[:, :3] *= np.linspace(0.4 if allegro else 0.0, 1.0, B, dtype=np.float32)[:, None]
    dpose, tf = fk_state(lib, be, hand, pack, pose)
    L = len(pack["mesh_frame"])
    x = be.from_host(sp[None], np.float32)
    dist = be.empty((B, n_pts), np.float32)
    e1, e2 = be.empty((B,), np.float32), be.empty((B,), np.float32)
    sc = be.empty((int(lib.gqp_hand_distance_scratch_floats(1, B, n_pts, 0)),), np.float32)
    g1, g2 = be.empty((B, L, 12), np.float32), be.empty((B, L, 12), np.float32)
    _lib.check(lib.gqp_hand_distance(hand, be.ptr(tf), be.ptr(x), 1, B, n_pts, 0, be.ptr(dist), be.ptr(e1), be.ptr(sc),
                                     100.0, be.ptr(g1), be.stream), lib)
    _lib.check(lib.gqp_hand_distance(hand, be.ptr(tf), be.ptr(x), 1, B, n_pts, 0, None, be.ptr(e2), be.ptr(sc), 100.0,
                                     be.ptr(g2), be.stream), lib)
    e1h, e2h, g1h, g2h, dh = [be.to_host(a) for a in (e1, e2, g1, g2, dist)]
    assert (e1h > 0).sum() >= (B // 4 if allegro else max(1, B // 8)), "poses do not penetrate"
    # same winners; the culled path sums in 2^-40 m fixed point, the full path in fp32 in a fixed order
    np.testing.assert_allclose(e2h, e1h, rtol=1e-6, atol=1e-9)
    # atomics: order differs; deep poses reach |g| ~ 1e3, the absolute slack follows the largest entry
    np.testing.assert_allclose(g2h, g1h, rtol=1e-4, atol=1e-4 * max(1.0, float(np.abs(g1h).max()) / 10))
    nb = min(B, 6)
    o = HandOracle(pack, torch.float32)
    o.set_parameters(torch.tensor(pose[:nb]), torch.tensor(idx[:nb]).long())
    want = o.cal_distance(torch.tensor(sp).unsqueeze(0).repeat(nb, 1, 1)).numpy()
    np.testing.assert_allclose(dh[:nb], want, atol=1e-5)
    np.testing.assert_allclose(e1h[:nb], np.maximum(want, 0).sum(-1), rtol=1e-4, atol=1e-5)
    lib.gqp_hand_destroy(hand)


def check_engine_vs_oracle(lib, be, B=8, n_pts=256, n_c=12, T=4, seed=3, hull=False, hand="allegro", max_excused=1):
    pack = hand_pack.load_pack(hand)
    v, f = synth.make_mesh("hull" if hull else "sphere", 400, seed)
    fv = synth.face_vertices(v, f)
    sp = synth.sample_surface(v, f, n_pts, seed=seed)
    pose, idx = synth.random_grasp_init(pack, [v], B, n_c, seed)
    pose[:, :3] *= 0.7
    eng = engine.FitEngine(lib, be, pack, [fv], sp[None], B, pose, idx.astype(np.int32))
    orc = FitLoopOracle(pack, [fv], [sp], B, pose, idx, dtype=torch.float64)
    H = be.to_host
    ex = lambda b: tol.normal_unstable(H(eng.cpts)[b], fv)
    tol.assert_energy(H(eng.energy), orc.energy.numpy(), f"{hand} n_c={n_c} initial energy")
    tol.assert_grad(eng.first_grad, orc.first_grad.numpy(), f"{hand} n_c={n_c} first gradient")
    rng = np.random.default_rng(0)
    n_acc = 0
    for t in range(T):
        us, ir, ua = rng.random((B, n_c), dtype=np.float32), rng.integers(0, len(pack["cand"]), (B, n_c)).astype(np.int32), rng.random(B, dtype=np.float32)
        eng.step(be.from_host(us), be.from_host(ir), be.from_host(ua))
        out = orc.iterate(torch.tensor(us, dtype=torch.float64), torch.tensor(ir).long(), torch.tensor(ua, dtype=torch.float64))
        np.testing.assert_array_equal(H(eng.idx_new), out["idx_prop"].numpy())
        tol.assert_energy(H(eng.new_energy), out["new_energy"].numpy(), f"{hand} n_c={n_c} step {t} new energy", excuse=ex,
                          max_excused=max_excused)
        tol.assert_grad(H(eng.grad_new), out["new_grad"].numpy(), f"{hand} n_c={n_c} step {t} gradient", excuse=ex,
                        max_excused=max_excused)
        np.testing.assert_array_equal(H(eng.accept).astype(bool), out["accept"].numpy())
        np.testing.assert_allclose(H(eng.energy), orc.energy.numpy(), rtol=tol.ENERGY_RTOL)
        np.testing.assert_allclose(H(eng.pose), orc.hand.hand_pose.detach().numpy(), rtol=1e-4, atol=1e-5)
        n_acc += int(out["accept"].sum())
    eng.close()
    return n_acc


def check_trajectory(lib, be, T=500, B=6, n_pts=128, n_c=12, seed=21, tie=2e-3, hand="allegro", fc_tol=1e-5):
    """A T-step fixed-noise trajectory of the engine, every step checked against the fp64 oracle started from the
    engine's own state (teacher forcing: two chaotic trajectories cannot be compared open-loop for 500 steps).
    Stated tolerances (tests/tol.py): proposal 1e-5 abs, new energy and gradient 1e-4 relative, contact indices
    bit-exact, accept decisions bit-exact except where the uniform draw lies within `tie` of the acceptance probability
    (must stay below 1 %).  -> (accepted, borderline, worst energy error)"""
    pack = hand_pack.load_pack(hand)
    v, f = synth.make_mesh("hull", 400, seed)
    fv = synth.face_vertices(v, f)
    sp = synth.sample_surface(v, f, n_pts, seed=seed)
    pose, idx = synth.random_grasp_init(pack, [v], B, n_c, seed)
    pose[:, :3] *= 0.75
    eng = engine.FitEngine(lib, be, pack, [fv], sp[None], B, pose, idx.astype(np.int32))
    orc = FitLoopOracle(pack, [fv], [sp], B, pose, idx, dtype=torch.float64)
    H = be.to_host
    rng = np.random.default_rng(seed)
    n_acc = n_tie = 0
    worst = worst_g = 0.0
    outliers, grad_events, grad_errs, grad_unexplained = [], 0, [], []
    E_TOL = tol.ENERGY_RTOL
    for t in range(T):
        orc.load_state(H(eng.pose), H(eng.idx), H(eng.grad), H(eng.energy), H(eng.ema), H(eng.step_count))
        us, ir, ua = rng.random((B, n_c), dtype=np.float32), rng.integers(0, len(pack["cand"]), (B, n_c)).astype(np.int32), rng.random(B, dtype=np.float32)
        eng.step(be.from_host(us), be.from_host(ir), be.from_host(ua))
        out = orc.iterate(torch.tensor(us, dtype=torch.float64), torch.tensor(ir).long(), torch.tensor(ua, dtype=torch.float64))
        np.testing.assert_array_equal(H(eng.idx_new), out["idx_prop"].numpy(), err_msg=f"step {t}")
        np.testing.assert_allclose(H(eng.pose_new), out["proposal"].numpy(), rtol=1e-5, atol=1e-5, err_msg=f"step {t}")
        e_new, want = H(eng.new_energy), out["new_energy"].numpy()
        rel = np.abs(e_new - want) / np.abs(want)
        flagged = rel > E_TOL
        if flagged.any():  # discontinuity events, see the assertion below
            cp64, nn64 = torch.tensor(H(eng.cpts), dtype=torch.float64), torch.tensor(H(eng.obj_nrm), dtype=torch.float64)
            E_in, _, info = oqp.graspqp_energy(cp64, nn64, orc.obj.cog)
            sv = torch.linalg.svdvals(info["F"])
            e_fc = H(eng.losses_new)[engine.TERMS.index("E_fc")]
        for b in np.nonzero(flagged)[0]:
            terms = {k: float(H(eng.losses_new)[i][b] - out["new_losses"][k][b]) for i, k in enumerate(engine.TERMS)}
            # ... and the engine's object normals are the reference kernel's, bit for bit, at the engine's own contacts
            _, sg_w, nr_w, _ = osdf.compute_sdf_port(H(eng.cpts)[b], fv)
            np.testing.assert_array_equal(H(eng.obj_nrm)[b], nr_w * sg_w[:, None].astype(np.float32), err_msg=f"step {t}")
            # E_fc of the oracle's QP + svd term on the ENGINE's own contacts / normals.  When the grasp matrix is rank
            # deficient to fp64 rounding (sigma_6 < 1e-12 sigma_1: all contacts on one line, e.g. duplicated contact
            # indices at n_c = 4) the reference's term prod(sigma)^(1/6) (span.py:402) is rounding noise to the power
            # 1/6 in ANY arithmetic; E_fc can then only agree up to svd_gain x that term.
            fc_err = abs(float(e_fc[b]) - E_in[b].item()) / abs(E_in[b].item())
            fc_allow = fc_tol if sv[b, -1] > 1e-12 * sv[b, 0] else fc_tol + 2 * 0.1 * float(info["svd"][b])
            outliers.append((t, int(b), float(rel[b]), terms, fc_err, fc_allow))
        worst = max(worst, float(np.max(np.where(flagged, 0.0, rel))))
        # gradient of the proposal: 1e-4 relative per pose column on every candidate that is not a conditioning event
        ge = tol.grad_error(H(eng.grad_new), out["new_grad"].numpy()).max(1)
        with torch.no_grad():
            _, n_orc = orc.obj.cal_distance(orc.hand.contact_points.detach())
        slack = tol.normal_slack(H(eng.obj_nrm), n_orc.numpy())
        for b in np.nonzero((ge > tol.GRAD_TOL) & ~flagged)[0]:
            # the gradient feels a rotated contact normal before the energy does: same evidence as for the energies
            if ge[b] <= tol.GRAD_TOL + slack[b] or tol.normal_unstable(H(eng.cpts)[b], fv):
                grad_events += 1
            else:
                grad_unexplained.append((t, int(b), float(ge[b])))
        grad_errs.append(ge[~flagged & (ge <= tol.GRAD_TOL)])
        p_acc = np.exp(np.minimum((orc_energy_before(out) - want) / out["temperature"].numpy(), 0.0))
        got, exp = H(eng.accept).astype(bool), out["accept"].numpy()
        border = (np.abs(ua - p_acc) < tie) | flagged
        assert np.array_equal(got[~border], exp[~border]), f"step {t}: accept mismatch away from a tie"
        n_tie += int((got != exp).sum())
        n_acc += int(got.sum())
    eng.close()
    assert n_tie <= max(1, T * B // 100), f"{n_tie} borderline accept mismatches"
    # Conditioning events of the reference's own formulation (found with tools/debug_traj.py): the object normal at a
    # contact is unit(p - c) * sign (object_model.py:246) and is not Lipschitz in p near nearest-face ties and on the
    # surface itself; the contact points of the two runs differ by one fp32 ulp (fp32 FK here, fp64 FK in the oracle),
    # which occasionally rotates ONE contact normal by 1e-4..1e-2 rad, and E_fc sees that.  Such steps are re-checked
    # with the reference SDF arithmetic on the engine's own fp32 contact points (normals must be bit-identical) and the
    # oracle's QP on the engine's own contacts/normals (must agree to fc_tol = 1e-5); a nearest-face tie can even flip a
    # normal's sign (one 9 % event seen in 3000 candidate-steps).  Allowed for at most 1 % of candidate-steps (0.1-0.3 %
    # observed); everything else is within 1e-4.
    for o in outliers:
        print("trajectory outlier (step, candidate, rel err, per-term diff, E_fc rel err on the engine's inputs, allowed):", o)
    ge = np.concatenate(grad_errs) if grad_errs else np.zeros(1)
    tol.MEASURED.append((f"{hand} n_c={n_c} {T}-step trajectory", "grad", float(ge.max()), float(np.median(ge))))
    print(f"[traj] {hand} n_c={n_c}: {T * B} candidate-steps, {len(outliers)} energy conditioning events, "
          f"{grad_events} gradient-only events with a nearest-face tie at a contact, {len(grad_unexplained)} without that "
          f"evidence {grad_unexplained[:5]}; elsewhere energy error max {worst:.2e}, gradient error max {ge.max():.2e} "
          f"median {np.median(ge):.2e}")
    assert len(outliers) <= max(1, T * B // 100), outliers
    # gradients: every excess needs the tie evidence (<= 5 % of candidate-steps on these coarse 400-face meshes, where a
    # contact 3 cm from the surface sits near a Voronoi boundary of the faces far more often than on a 10k-face mesh:
    # 0-1 of 16 rows in the benched-shape checks); unexplained excesses: at most 0.2 %
    assert grad_events <= max(1, T * B // 20), f"{grad_events} candidate-steps with a gradient error above {tol.GRAD_TOL}"
    assert len(grad_unexplained) <= max(1, T * B // 500), grad_unexplained
    assert all(o[4] < o[5] and abs(o[3]["E_pen"]) < 1e-5 for o in outliers), outliers
    return n_acc, n_tie, worst


def orc_energy_before(out):
    return out["energy_before"].numpy()


def link_grid(lib, hand, link):
    geom, dim = np.zeros(10, np.float32), np.zeros(3, np.int32)
    _lib.check(lib.gqp_hand_link_grid(hand, link, geom.ctypes.data, dim.ctypes.data, None, 0), lib)
    bits = None
    if dim.prod() > 0:
        words = (int(dim.prod()) + 31) // 32
        bits = np.zeros(words, np.uint32)
        _lib.check(lib.gqp_hand_link_grid(hand, link, geom.ctypes.data, dim.ctypes.data, bits.ctypes.data, words), lib)
    return geom, dim, bits


def link_planes(lib, hand, link):
    n = lib.gqp_hand_link_planes(hand, link, None, 0, None)
    assert n >= 0, lib.gqp_last_error()
    slab = np.zeros(3, np.float32)
    lib.gqp_hand_link_planes(hand, link, None, 0, slab.ctypes.data)
    n_a = int(slab[2])
    table = np.zeros((max(n + n_a, 1), 4), np.float32)
    assert lib.gqp_hand_link_planes(hand, link, table.ctypes.data, n + n_a, slab.ctypes.data) == n
    return table[:n], slab[:2], table[n:n + n_a, :3]


def in_tie_slab(q, planes, slab, lo, hi, anchors=None):
    """csrc/gqp_pen.cu:pen_in_tie_slab for points (rad = 0), in fp32 like the kernel: the coarse slab (distance bounded by
    the link's bounding sphere), then the slab with the distance bounded by the nearest anchor vertex"""
    q = np.asarray(q, np.float32)
    if len(planes) == 0:
        return np.zeros(len(q), bool)
    ctr, half = np.float32(0.5) * (lo + hi), np.float32(0.5) * (hi - lo)
    D = np.sqrt(((q - ctr) ** 2).sum(1, dtype=np.float32)) + np.sqrt((half ** 2).sum(dtype=np.float32))
    hgt = q @ planes[:, :3].T - planes[:, 3]
    coarse = ((hgt <= slab[1] * D[:, None]) & (hgt >= -slab[0] * D[:, None])).any(1)
    if anchors is None or len(anchors) == 0:
        return coarse
    d2 = np.minimum(D * D, ((q[:, None, :] - anchors[None]) ** 2).sum(-1, dtype=np.float32).min(1))
    D2 = np.sqrt(d2) * np.float32(1.001)
    fine = ((hgt <= slab[1] * D2[:, None]) & (hgt >= -slab[0] * D2[:, None])).any(1)
    return coarse & fine


def check_culling_proof(lib, sign_fn, hand_name, per_cell=4, n_far=20000, seed=0, max_cells=None):
    """Direct test of the culling proof that carries > 98 % of H4's work (csrc/gqp_bvh.cpp: gqp_mesh_is_nice +
    gqp_outside_grid_build): for every link with an outside grid, for every SET cell, points drawn inside the cell
    (uniform + the eight corners pulled in by 1e-6 of the cell) must get sign +1 from the reference's brute-force
    arithmetic (`sign_fn(points, face_verts) -> int32 signs`: the TorchSDF port on the CPU, gqp_sdf_forward on the GPU);
    likewise points outside the link's box out to 1.5 m.  The cell of a point is computed exactly as pen_outside() does
    (fp32 (q - origin) * inv_cell, truncation).  -> (links with a grid, cells checked, points checked)"""
    pack = hand_pack.load_pack(hand_name)
    hand = hand_pack.create_hand(lib, pack)
    rng = np.random.default_rng(seed)
    L = len(pack["mesh_frame"])
    n_links = n_cells = n_points = 0
    slab_seen = [0, 0]
    for l in range(L):
        f0, f1 = int(pack["face_offset"][l]), int(pack["face_offset"][l + 1])
        geom, dim, bits = link_grid(lib, hand, l)
        if bits is None or f1 == f0:
            continue
        fv = np.ascontiguousarray(pack["face_verts"][f0:f1], np.float32).reshape(-1, 3, 3)
        org, inv_cell, lo, hi = geom[:3], geom[3], geom[4:7], geom[7:10]
        cell = np.float32(1.0) / inv_cell
        idx = np.nonzero(np.unpackbits(bits.view(np.uint8), bitorder="little")[:int(dim.prod())])[0]
        if max_cells is not None and len(idx) > max_cells:
            idx = rng.choice(idx, max_cells, replace=False)
        cx, cy, cz = idx % dim[0], (idx // dim[0]) % dim[1], idx // (dim[0] * dim[1])
        c0 = np.stack([cx, cy, cz], 1).astype(np.float64)
        corners = np.array([[i, j, k] for i in (1e-6, 1 - 1e-6) for j in (1e-6, 1 - 1e-6) for k in (1e-6, 1 - 1e-6)])
        frac = np.concatenate([rng.random((len(idx), per_cell, 3)), np.broadcast_to(corners, (len(idx), 8, 3))], 1)
        pts = (org.astype(np.float64) + (c0[:, None, :] + frac) * float(cell)).reshape(-1, 3).astype(np.float32)
        # keep the points the KERNEL would file under a set cell (fp32 arithmetic of pen_outside)
        fcell = (pts - org) * inv_cell
        ic = fcell.astype(np.int32)
        ok = (fcell >= 0).all(1) & (ic < dim).all(1)
        lin = (ic[:, 2].astype(np.int64) * dim[1] + ic[:, 1]) * dim[0] + ic[:, 0]
        lin = np.where(ok, lin, 0)
        setbit = ((bits[lin >> 5] >> (lin & 31).astype(np.uint32)) & 1).astype(bool) & ok
        inside_box = ((pts >= lo) & (pts <= hi)).all(1)
        pts = pts[setbit & inside_box]
        # far field: outside the link's box (pen_outside returns true before looking at the grid)
        d = rng.normal(size=(n_far, 3))
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        ctr, half = 0.5 * (lo + hi), 0.5 * (hi - lo)
        far = (ctr + d * (np.linalg.norm(half) * (1.0 + rng.random((n_far, 1)) ** 3 * 15.0) + 1e-4)).astype(np.float32)
        far = far[~((far >= lo) & (far <= hi)).all(1)]
        near = (ctr + (rng.random((n_far, 3)) * 2 - 1) * half * 1.3).astype(np.float32)  # just outside the box faces
        near = near[~((near >= lo) & (near <= hi)).all(1)]
        extra = [pts, far, near]
        planes, slab, anchors = link_planes(lib, hand, l)
        if len(planes):
            # stress the slab's boundary: points thrown at the planes, heights within +-3e-3 of their distance bound
            m = n_far // 2
            base = (ctr + (rng.random((m, 3)) * 2 - 1) * (half + 0.4)).astype(np.float64)
            k = rng.integers(0, len(planes), m)
            nk, ak = planes[k, :3].astype(np.float64), planes[k, 3].astype(np.float64)
            Dk = np.linalg.norm(base - ctr, axis=1) + np.linalg.norm(half)
            newh = (rng.random(m) * 3.1e-3 - 3e-3) * Dk
            base = (base + (newh - ((base * nk).sum(1) - ak))[:, None] * nk).astype(np.float32)
            extra.append(base[~((base >= lo) & (base <= hi)).all(1)])
        q = np.concatenate(extra)
        # the tie slabs behind the planes of faces at right-angle creases are never culled (gqp_mesh_tie_planes): there
        # the reference's fp32 first-minimum rule CAN say -1 for an outside point; count those as evidence
        slab_pts = in_tie_slab(q, planes, slab, lo, hi, anchors)
        n_slab_neg = 0
        if slab_pts.any():
            n_slab_neg = int((sign_fn(q[slab_pts], fv) != 1).sum())
        slab_seen[0] += int(slab_pts.sum()); slab_seen[1] += n_slab_neg
        q = q[~slab_pts]
        sg = sign_fn(q, fv)
        bad = np.nonzero(sg != 1)[0]
        assert len(bad) == 0, (f"{hand_name} link {l}: {len(bad)} of {len(q)} points in culled regions get sign -1 from "
                               f"the reference arithmetic, e.g. {q[bad[:3]]} (first {len(pts)} are grid points)")
        n_links += 1
        n_cells += len(idx)
        n_points += len(q)
    lib.gqp_hand_destroy(hand)
    print(f"[culling proof] {hand_name}: {n_links} links with a grid, {n_cells} set cells, {n_points} points, all +1; "
          f"{slab_seen[0]} further points fell into never-culled tie slabs, {slab_seen[1]} of them with sign -1")
    return n_links, n_cells, n_points


def check_benched_state(lib, be, name="config2", iters=(0, 50, 500), per_object=2, seed=0, n_objects=None,
                        batch_each=None):
    """Parity AT THE BENCHED SHAPES: builds the workload bench.py times (full candidate count, 2500 surface samples,
    10k / 50k-face meshes), runs the engine to iteration n (Philox noise), then takes `per_object` random candidates of
    every object OF THAT STATE and checks one more iteration (injected noise, shared by both sides) against the fp64
    oracle holding only those candidates, seeded with the batch-wide statistics of the full batch: (g^2).mean(0)
    (optimizer.py:231), the per-object z-scores (fit.py:403-406) and qpth's batch-wide exit iteration.
    Tolerances: tests/tol.py (1e-4 energies / gradients, nearest-face-tie evidence for excused rows), indices bit-exact,
    accept decisions bit-exact away from ties.  -> dict of measured errors per checkpoint"""
    from graspqp_b200 import workload
    w = workload.build(name, n_objects=n_objects, batch_each=batch_each)
    pack, B, n_c, O = w["pack"], w["batch_each"], w["n_contact"], len(w["face_verts"])
    Bt = O * B
    eng = engine.FitEngine(lib, be, pack, w["face_verts"], w["surface"], B, w["pose"], w["idx"])
    H = be.to_host
    rng = np.random.default_rng(seed)
    sub = np.concatenate([o * B + np.sort(rng.choice(B, per_object, replace=False)) for o in range(O)])
    n_cand = len(pack["cand"])
    report = {}
    done = 0
    orc = None
    for n in iters:
        while done < n:
            eng.step()
            done += 1
        st = {k: H(getattr(eng, k)) for k in ("pose", "idx", "grad", "energy", "ema", "step_count")}
        us, ir, ua = rng.random((Bt, n_c), dtype=np.float32), rng.integers(0, n_cand, (Bt, n_c)).astype(np.int32), rng.random(Bt, dtype=np.float32)
        eng.step(be.from_host(us), be.from_host(ir), be.from_host(ua))
        done += 1
        if orc is None:
            orc = FitLoopOracle(pack, w["face_verts"], list(w["surface"]), per_object, st["pose"][sub], st["idx"][sub],
                                dtype=torch.float64)
        orc.load_state(st["pose"][sub], st["idx"][sub], st["grad"][sub], st["energy"][sub], st["ema"][sub],
                       st["step_count"][sub])
        t64 = lambda a: torch.tensor(a, dtype=torch.float64)
        out = orc.iterate(t64(us[sub]), torch.tensor(ir[sub]).long(), t64(ua[sub]), grad2_mean=H(eng.grad2_mean),
                          z=H(eng.z)[sub], qp_stop=int(H(eng.qp_stop)[0]))
        cp = H(eng.cpts)
        cp_orc = orc.hand.contact_points.detach().numpy()
        ex = lambda r: (tol.normals_differ(cp[sub[r]], cp_orc[r], w["face_verts"][sub[r] // B]) or
                        tol.normal_unstable(cp[sub[r]], w["face_verts"][sub[r] // B]))
        lab = f"{name} Bt={Bt} iteration {n}"
        np.testing.assert_array_equal(H(eng.idx_new)[sub], out["idx_prop"].numpy(), err_msg=lab)
        np.testing.assert_allclose(H(eng.pose_new)[sub], out["proposal"].numpy(), rtol=1e-5, atol=1e-5, err_msg=lab)
        e_probe = np.abs(H(eng.new_energy)[sub] - out["new_energy"].numpy()) / np.abs(out["new_energy"].numpy())
        for r in np.nonzero(e_probe > tol.ENERGY_RTOL)[0]:  # diagnostics before the assertion
            print(f"[benched] {lab}: row {r} (candidate {sub[r]}) energy {H(eng.new_energy)[sub[r]]:.6g} vs oracle "
                  f"{out['new_energy'][r].item():.6g}; per term (engine, oracle): " +
                  ", ".join(f"{t}=({H(eng.losses_new)[k][sub[r]]:.6g}, {out['new_losses'][t][r].item():.6g})"
                            for k, t in enumerate(engine.TERMS)))
        e_rel = tol.assert_energy(H(eng.new_energy)[sub], out["new_energy"].numpy(), lab + " new energy", excuse=ex,
                                  max_excused=max(1, len(sub) // 8))
        excused = set(int(r) for r in np.nonzero(e_rel > tol.ENERGY_RTOL)[0])
        e_tot = np.abs(out["new_energy"].numpy())
        for k, term in enumerate(engine.TERMS):
            # every term within 1e-4 of itself, or (a small term next to a large one) its weighted error within 1e-4 of
            # the candidate's total energy, which is what the north star's tolerance is stated on
            got, want = H(eng.losses_new)[k][sub], out["new_losses"][term].numpy()
            keep = np.array([r not in excused or term in ("E_pen", "E_spen", "E_joints") for r in range(len(sub))])
            err = np.abs(got - want)
            ok = (err <= 1e-4 * np.abs(want) + 1e-6) | (engine.DEFAULT_WEIGHTS[term] * err <= 1e-4 * e_tot)
            if term in ("E_fc", "E_dis"):
                # these two read the object normal unit(p - c): a row whose two evaluations do not see the same normals (the
                # evidence functions of tests/tol.py) is a conditioning event of the reference's formulation for THIS term
                # even when its total energy still agrees; each such row needs the evidence, at most len/8 of them
                cond = [r for r in np.nonzero(keep & ~ok)[0] if ex(r)]
                for r in cond:
                    print(f"[benched] {lab} {term}: row {r} {got[r]:.6g} vs {want[r]:.6g} — object normals of the two "
                          "evaluations differ at one of its contacts (excused)")
                assert len(cond) <= max(1, len(sub) // 8), f"{lab} {term}: {len(cond)} rows with differing normals"
                ok[cond] = True
            assert ok[keep].all(), f"{lab} {term}: {got[keep & ~ok]} vs {want[keep & ~ok]}"
        ge_probe = tol.grad_error(H(eng.grad_new)[sub], out["new_grad"].numpy())
        for r in np.nonzero(ge_probe.max(1) > tol.GRAD_TOL)[0]:  # diagnostics before the assertion
            Fm = oqp.grasp_matrix(orc.hand.contact_points.detach()[r:r + 1], torch.tensor(H(eng.obj_nrm)[sub[r:r + 1]],
                                  dtype=torch.float64), orc.obj.cog[r:r + 1])
            sv = torch.linalg.svdvals(Fm)[0]
            with torch.no_grad():
                dl = orc.hand.cal_distance(orc.obj.surface_points_tensor, per_link=True)[:, r]  # (links, samples)
            top2 = torch.topk(dl, 2, dim=0)[0]
            print(f"[benched] {lab}: row {r} gradient error {ge_probe[r].max():.2e} in column {int(ge_probe[r].argmax())}; "
                  f"sigma_min/sigma_max of F {float(sv[-1] / sv[0]):.2e}; min |object distance| at a contact "
                  f"{float(np.abs(H(eng.obj_dist)[sub[r]]).min()):.2e}; samples with |depth| < 2e-6: "
                  f"{int((top2[0].abs() < 2e-6).sum())}; penetrating samples with a link tie < 2e-6: "
                  f"{int(((top2[0] > 0) & ((top2[0] - top2[1]) < 2e-6)).sum())}; tie evidence {ex(int(r))}")
        with torch.no_grad():
            _, n_orc = orc.obj.cal_distance(orc.hand.contact_points.detach())
        slack = tol.normal_slack(H(eng.obj_nrm)[sub], n_orc.numpy())
        g_rows = tol.assert_grad(H(eng.grad_new)[sub], out["new_grad"].numpy(), lab + " gradient", excuse=ex,
                                 max_excused=max(1, len(sub) // 8), slack=slack)
        p_acc = np.exp(np.minimum((out["energy_before"].numpy() - out["new_energy"].numpy()) / out["temperature"].numpy(), 0.0))
        got, exp = H(eng.accept)[sub].astype(bool), out["accept"].numpy()
        border = (np.abs(ua[sub] - p_acc) < 2e-3) | (e_rel > tol.ENERGY_RTOL)
        assert np.array_equal(got[~border], exp[~border]), f"{lab}: accept mismatch away from a tie"
        report[n] = dict(energy_max=float(np.where(e_rel > tol.ENERGY_RTOL, 0, e_rel).max()),
                         grad_max=float(np.where(g_rows > tol.GRAD_TOL, 0, g_rows).max()), excused=len(excused),
                         qp_stop=int(H(eng.qp_stop)[0]), accepted=int(got.sum()),
                         e_pen_mean=float(H(eng.losses)[engine.TERMS.index("E_pen")].mean()))
        print(f"[benched] {lab}: {report[n]}")
    eng.close()
    return report


def check_optional_terms(lib, be, B=4, n_pts=64, n_c=12, T=2, seed=3, hand="allegro", w_prior=5.0, w_wall=50.0):
    """(f3) E_prior / E_wall (core/energy.py:68-77) switched on by their weights: engine vs fp64 oracle, per-term energies,
    total energy, gradient, accept decisions.  Both sides read the same hand surface samples.  -> (mean E_prior, mean E_wall)"""
    pack = hand_pack.load_pack(hand)
    v, f = synth.make_mesh("sphere", 200, seed)
    fv = synth.face_vertices(v, f)
    sp = synth.sample_surface(v, f, n_pts, seed=seed)
    pose, idx = synth.random_grasp_init(pack, [v], B, n_c, seed)
    pose[:, :3] *= 0.7
    w = dict(E_prior=w_prior, E_wall=w_wall)
    surf = hand_pack.surface_points(pack, 128)
    eng = engine.FitEngine(lib, be, pack, [fv], sp[None], B, pose, idx.astype(np.int32), weights=w, hand_surface=surf)
    orc = FitLoopOracle(pack, [fv], [sp], B, pose, idx, dtype=torch.float64, weights=w, hand_surface=surf)
    H = be.to_host
    assert eng.terms == engine.TERMS + tuple(k for k in engine.OPTIONAL_TERMS if w[k] > 0)
    tol.assert_energy(H(eng.energy), orc.energy.numpy(), f"{hand} optional terms: initial energy")
    for k, name in enumerate(eng.terms):
        np.testing.assert_allclose(H(eng.losses)[k], orc.losses[name].numpy(), rtol=1e-4, atol=1e-6, err_msg=name)
    tol.assert_grad(eng.first_grad, orc.first_grad.numpy(), f"{hand} optional terms: first gradient")
    ex = lambda b: tol.normal_unstable(H(eng.cpts)[b], fv)
    rng = np.random.default_rng(0)
    for t in range(T):
        us, ir, ua = rng.random((B, n_c), dtype=np.float32), rng.integers(0, len(pack["cand"]), (B, n_c)).astype(np.int32), rng.random(B, dtype=np.float32)
        eng.step(be.from_host(us), be.from_host(ir), be.from_host(ua))
        out = orc.iterate(torch.tensor(us, dtype=torch.float64), torch.tensor(ir).long(), torch.tensor(ua, dtype=torch.float64))
        for k, name in enumerate(eng.terms):
            np.testing.assert_allclose(H(eng.losses_new)[k], out["new_losses"][name].numpy(), rtol=1e-4, atol=1e-6, err_msg=name)
        tol.assert_energy(H(eng.new_energy), out["new_energy"].numpy(), f"{hand} optional terms: step {t} new energy", excuse=ex)
        tol.assert_grad(H(eng.grad_new), out["new_grad"].numpy(), f"{hand} optional terms: step {t} gradient", excuse=ex)
        np.testing.assert_array_equal(H(eng.accept).astype(bool), out["accept"].numpy())
    ep = float(H(eng.losses_new)[eng.terms.index("E_prior")].mean()) if w_prior > 0 else 0.0
    ew = float(H(eng.losses_new)[eng.terms.index("E_wall")].mean()) if w_wall > 0 else 0.0
    eng.close()
    return ep, ew
