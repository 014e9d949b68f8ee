// M1/M2/M4 — MALA* proposal, accept/reject and the caller's bookkeeping.
//   MalaStar.try_step    core/optimizer.py:199-273      MalaStar.accept_step  core/optimizer.py:289-340
//   z-score, weighted sum, masked commit                scripts/fit.py:403-406, 434-458
// The reference runs ~15 element-wise launches and three host syncs per iteration here; these are three launches, no
// sync.  State rows are candidates; everything is HBM-streaming over a few KB per candidate (bound: launch latency).
// Randomness: injected arrays (parity runs) or Philox4x32-10 keyed by (seed, iteration, row) (production runs).
#include "gqp_philox.cuh"
#include "graspqp_b200.h"

#define STATS_ROWS 64

// blocks [0, n_gblk): partial sums of grad^2 over 64 rows; blocks [n_gblk, n_gblk + n_obj): z-score of one object
__global__ void __launch_bounds__(256)
mala_stats_kernel(const float *__restrict__ grad, const float *__restrict__ energy, int n_obj, int batch_each, int D,
                  int n_gblk, int clip_grad, float *__restrict__ partial, float *__restrict__ z) {
    const int Bt = n_obj * batch_each;
    if ((int)blockIdx.x < n_gblk) {
        const int r0 = blockIdx.x * STATS_ROWS, r1 = min(Bt, r0 + STATS_ROWS);
        for (int d = threadIdx.x; d < D; d += blockDim.x) {
            float s = 0.f;
            for (int r = r0; r < r1; ++r) {
                float g = grad[(size_t)r * D + d];
                if (clip_grad) g = (g != g) ? 0.f : fminf(fmaxf(g, -100.f), 100.f);  // :211-213 (clip keeps NaN, then NaN -> 0)
                s += g * g;
            }
            partial[(size_t)blockIdx.x * D + d] = s;
        }
        return;
    }
    // z-score (fit.py:403-406): mean and UNBIASED std over the object's candidates, fixed-order tree reduction
    __shared__ float red[256];
    const int o = blockIdx.x - n_gblk;
    const float *e = energy + (size_t)o * batch_each;
    float s = 0.f;
    for (int i = threadIdx.x; i < batch_each; i += blockDim.x) s += e[i];
    red[threadIdx.x] = s;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if ((int)threadIdx.x < w) red[threadIdx.x] += red[threadIdx.x + w];
        __syncthreads();
    }
    const float mean = red[0] / (float)batch_each;
    __syncthreads();
    s = 0.f;
    for (int i = threadIdx.x; i < batch_each; i += blockDim.x) {
        float dlt = e[i] - mean;
        s += dlt * dlt;
    }
    red[threadIdx.x] = s;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if ((int)threadIdx.x < w) red[threadIdx.x] += red[threadIdx.x + w];
        __syncthreads();
    }
    const float sd = sqrtf(red[0] / (float)(batch_each - 1));
    for (int i = threadIdx.x; i < batch_each; i += blockDim.x) z[(size_t)o * batch_each + i] = (e[i] - mean) / sd;
}

__global__ void mala_stats_final_kernel(const float *__restrict__ partial, int n_gblk, int D, int Bt,
                                        float *__restrict__ grad2_mean) {
    int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= D) return;
    float s = 0.f;
    for (int k = 0; k < n_gblk; ++k) s += partial[(size_t)k * D + d];
    grad2_mean[d] = s / (float)Bt;
}

// optimizer.py:199-273
__global__ void __launch_bounds__(128)
mala_propose_kernel(GqpMalaParams P, const float *__restrict__ pose, const float *__restrict__ grad,
                    const float *__restrict__ grad2_mean, float *__restrict__ ema, int32_t *__restrict__ step,
                    const int32_t *__restrict__ idx, int Bt, int D, int n_c, int n_cand,
                    const float *__restrict__ u_switch, const int32_t *__restrict__ idx_rand, uint64_t seed,
                    uint64_t iteration, float *__restrict__ pose_new, int32_t *__restrict__ idx_new,
                    float *__restrict__ step_size_out) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= Bt) return;
    const int st = step[b];
    const float s = P.step_size * powf(P.temperature_decay, (float)(st / P.stepsize_period));  // :209
    bool row_nan = false;
    for (int d = 0; d < D; ++d) {
        const size_t i = (size_t)b * D + d;
        float g = grad[i];
        if (P.clip_grad) g = (g != g) ? 0.f : fminf(fmaxf(g, -100.f), 100.f);  // :211-213
        float e = P.mu * grad2_mean[d] + (1.0f - P.mu) * ema[i];  // :233
        if (e != e) e = 0.f;                                       // :235-236
        ema[i] = e;
        const float pn = pose[i] - s * g / (sqrtf(e) + 1e-6f);  // :240
        row_nan |= pn != pn;
        pose_new[i] = pn;
    }
    if (row_nan)  // :242-244: a row with any NaN entry is zeroed as a whole
        for (int d = 0; d < D; ++d) pose_new[(size_t)b * D + d] = 0.f;
    for (int c = 0; c < n_c; ++c) {  // :252-257
        const size_t i = (size_t)b * n_c + c;
        float u;
        int32_t r;
        if (u_switch) {
            u = u_switch[i];
            r = idx_rand[i];
        } else {
            Philox p = philox4x32((uint32_t)b, (uint32_t)c, (uint32_t)iteration, (uint32_t)(iteration >> 32),
                                  (uint32_t)seed, (uint32_t)(seed >> 32));
            u = u01(p.c[0]);
            r = (int32_t)(((uint64_t)p.c[1] * (uint64_t)n_cand) >> 32);
        }
        idx_new[i] = (u < P.switch_possibility) ? r : idx[i];
    }
    step[b] = st + 1;  // :271
    if (step_size_out) step_size_out[b] = s;
}

// optimizer.py:289-340 + fit.py:434-458.  Accepted rows adopt the proposal; rejected rows keep pose / idx / grad.
__global__ void __launch_bounds__(128)
mala_accept_kernel(GqpMalaParams P, int Bt, int D, int n_c, int n_terms, const float *__restrict__ weights,
                   const float *__restrict__ losses_new, const float *__restrict__ z,
                   const uint8_t *__restrict__ reset_mask, const float *__restrict__ u_accept, uint64_t seed,
                   uint64_t iteration, const int32_t *__restrict__ step, int step_scalar,
                   const float *__restrict__ pose_new,
                   const int32_t *__restrict__ idx_new, const float *__restrict__ grad_new, float *__restrict__ pose,
                   int32_t *__restrict__ idx, float *__restrict__ grad, float *__restrict__ energy,
                   float *__restrict__ losses, float *__restrict__ new_energy_out, uint8_t *__restrict__ accept_out,
                   float *__restrict__ temperature_out, int reject_accumulates) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= Bt) return;
    float en = 0.f;
    for (int k = 0; k < n_terms; ++k) en += weights[k] * losses_new[(size_t)k * Bt + b];  // fit.py:434-438
    // per-row step counters (MalaStar, :301-303) or one counter for the batch (AnnealingDexGraspNet, :119-121)
    const int stp = step ? step[b] : step_scalar;
    float T = P.starting_temperature * powf(P.temperature_decay, (float)(stp / P.annealing_period));
    if (z) T = T * (1.0f + 0.5f * (1.0f + erff(z[b] * 0.70710678118654752f)));                            // :307-309
    float u;
    if (u_accept) u = u_accept[b];
    else {
        Philox p = philox4x32((uint32_t)b, 0xffffffffu, (uint32_t)iteration, (uint32_t)(iteration >> 32),
                              (uint32_t)seed, (uint32_t)(seed >> 32));
        u = u01(p.c[0]);
    }
    bool acc = u < expf((energy[b] - en) / T);  // :313
    if (reset_mask && reset_mask[b]) acc = true;  // :315-316
    if (acc) {
        for (int d = 0; d < D; ++d) {
            pose[(size_t)b * D + d] = pose_new[(size_t)b * D + d];
            grad[(size_t)b * D + d] = grad_new[(size_t)b * D + d];
        }
        for (int c = 0; c < n_c; ++c) idx[(size_t)b * n_c + c] = idx_new[(size_t)b * n_c + c];
        energy[b] = en;  // fit.py:454-458
        for (int k = 0; k < n_terms; ++k) losses[(size_t)k * Bt + b] = losses_new[(size_t)k * Bt + b];
    } else if (reject_accumulates) {
        // the iteration after a reset iteration: the reference hands a rejected row old + new gradient (see the header)
        for (int d = 0; d < D; ++d) grad[(size_t)b * D + d] += grad_new[(size_t)b * D + d];
    }
    if (new_energy_out) new_energy_out[b] = en;
    if (accept_out) accept_out[b] = acc ? 1 : 0;
    if (temperature_out) temperature_out[b] = T;
}

// M3: MalaStar.reset_envs (optimizer.py:275-287) + HandModel.set_parameters(..., env_mask) (hand_model.py:845-850,
// 816-831) as scripts/fit.py:408-422 applies them between try_step and the energy evaluation.
__global__ void __launch_bounds__(128)
mala_reset_kernel(int Bt, int D, int n_c, const float *__restrict__ z, float z_threshold,
                  const uint8_t *__restrict__ mask_in, const float *__restrict__ pose_init,
                  const int32_t *__restrict__ idx_init, float *__restrict__ pose_new, int32_t *__restrict__ idx_new,
                  int32_t *__restrict__ idx_eval, int32_t *__restrict__ step, float *__restrict__ ema,
                  uint8_t *__restrict__ mask_out, int32_t *__restrict__ count_out) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    bool m = false;
    if (b < Bt) {
        m = mask_in ? mask_in[b] != 0 : z[b] > z_threshold;  // fit.py:409 (NaN z: false, as torch's comparison)
        if (m) {
            for (int d = 0; d < D; ++d) {
                pose_new[(size_t)b * D + d] = pose_init[(size_t)b * D + d];
                ema[(size_t)b * D + d] = 0.f;  // :277
            }
            step[b] = 0;  // :276
        }
        for (int c = 0; c < n_c; ++c) {
            const int32_t r = idx_init[(size_t)b * n_c + c];
            if (m) idx_new[(size_t)b * n_c + c] = r;
            // hand_model.py:830-831 gathers the contact points of EVERY row with the freshly drawn indices, masked or
            // not, while contact_point_indices keeps the proposal's for unmasked rows: the energies of a reset
            // iteration are evaluated at idx_eval, the state that is committed is idx_new
            if (idx_eval) idx_eval[(size_t)b * n_c + c] = r;
        }
        mask_out[b] = m ? 1 : 0;
    }
    if (count_out) {
        const unsigned bal = __ballot_sync(GQP_FULL, m);
        if ((threadIdx.x & 31) == 0 && bal) atomicAdd(count_out, __popc(bal));
    }
}

extern "C" int gqp_mala_reset(int Bt, int D, int n_c, const float *z, float z_threshold, const uint8_t *mask_in,
                              const float *pose_init, const int32_t *idx_init, float *pose_new, int32_t *idx_new,
                              int32_t *idx_eval, int32_t *step, float *ema, uint8_t *mask_out, int32_t *count_out,
                              void *stream) {
    if (Bt < 0 || D <= 0 || n_c <= 0) return gqp_set_error(GQP_ERR_ARG, "gqp_mala_reset: bad argument");
    if (Bt == 0) return GQP_OK;
    if ((!z && !mask_in) || !pose_init || !idx_init || !pose_new || !idx_new || !step || !ema || !mask_out)
        return gqp_set_error(GQP_ERR_ARG, "gqp_mala_reset: null pointer");
    if (count_out) {
#ifndef GQP_EMU
        if (cudaMemsetAsync(count_out, 0, sizeof(int32_t), (cudaStream_t)stream) != cudaSuccess)
            return gqp_set_error(GQP_ERR_CUDA, "gqp_mala_reset: memset failed");
#else
        *count_out = 0;
#endif
    }
    GQP_LAUNCH(mala_reset_kernel, (Bt + 127) / 128, 128, 0, (cudaStream_t)stream, Bt, D, n_c, z, z_threshold, mask_in,
               pose_init, idx_init, pose_new, idx_new, idx_eval, step, ema, mask_out, count_out);
    GQP_CHECK_LAUNCH("gqp_mala_reset");
    return GQP_OK;
}

extern "C" int gqp_mala_stats(const float *grad, const float *energy, int n_obj, int batch_each, int D, int clip_grad,
                              float *grad2_mean, float *z, float *scratch, void *stream) {
    if (n_obj <= 0 || batch_each <= 0 || D <= 0 || !grad || !energy || !grad2_mean || !z || !scratch)
        return gqp_set_error(GQP_ERR_ARG, "gqp_mala_stats: bad argument");
    const int Bt = n_obj * batch_each, n_gblk = (Bt + STATS_ROWS - 1) / STATS_ROWS;
    GQP_LAUNCH(mala_stats_kernel, n_gblk + n_obj, 256, 0, (cudaStream_t)stream, grad, energy, n_obj, batch_each, D,
               n_gblk, clip_grad, scratch, z);
    GQP_CHECK_LAUNCH("gqp_mala_stats");
    GQP_LAUNCH(mala_stats_final_kernel, (D + 63) / 64, 64, 0, (cudaStream_t)stream, scratch, n_gblk, D, Bt,
               grad2_mean);
    GQP_CHECK_LAUNCH("gqp_mala_stats/final");
    return GQP_OK;
}

// ---- allocation-free variant (no counterpart in the reference) ----------------
// Callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc)
// cannot provide gqp_mala_stats' scratch.  This variant keeps it in a file-scope device array: every GPU that loads the
// library holds GQP_STATS_STATIC_FLOATS * 4 bytes = 256 KB until exit.  All calls share the array, which keeps the last
// call's contents: nothing in it has to start at zero (every partial sum the final pass reads is written by the first pass
// of the same call), but two calls that overlap on the GPU (different streams) race on it.
#define GQP_STATS_STATIC_FLOATS 65536
static __device__ __align__(256) float g_stats_static[GQP_STATS_STATIC_FLOATS];
#define GQP_MAX_DEVICES 64

// device address of the array for the CURRENT device, looked up once per device (host code must not pass the symbol itself:
// there its name is a host-side stand-in, not a device address)
static int stats_static_ptr(float **out) {
#ifndef GQP_EMU
    static float *cache[GQP_MAX_DEVICES] = {nullptr};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return gqp_set_error(GQP_ERR_CUDA, "gqp_mala_stats_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= GQP_MAX_DEVICES) return gqp_set_error(GQP_ERR_LIMIT, "gqp_mala_stats_noalloc: device %d out of range", dev);
    if (!cache[dev]) {
        void *ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, g_stats_static);
        if (e != cudaSuccess || !ptr)
            return gqp_set_error(GQP_ERR_CUDA, "gqp_mala_stats_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        cache[dev] = static_cast<float *>(ptr);
    }
    *out = cache[dev];
#else
    *out = g_stats_static;
#endif
    return GQP_OK;
}

extern "C" uint64_t gqp_mala_stats_noalloc_max_floats(void) { return GQP_STATS_STATIC_FLOATS; }

extern "C" int gqp_mala_stats_noalloc(const float *grad, const float *energy, int n_obj, int batch_each, int D,
                                      int clip_grad, float *grad2_mean, float *z, void *stream) {
    if (n_obj <= 0 || batch_each <= 0 || D <= 0 || !grad || !energy || !grad2_mean || !z)
        return gqp_set_error(GQP_ERR_ARG, "gqp_mala_stats_noalloc: bad argument");
    const long long Bt = (long long)n_obj * batch_each, need = (Bt + STATS_ROWS - 1) / STATS_ROWS * D;
    if (need > GQP_STATS_STATIC_FLOATS)
        return gqp_set_error(GQP_ERR_LIMIT,
                             "gqp_mala_stats_noalloc: %lld candidates x %d pose columns need %lld scratch floats, the static "
                             "array holds %d (use gqp_mala_stats with caller-owned scratch)", Bt, D, need, GQP_STATS_STATIC_FLOATS);
    float *scratch = nullptr;
    if (int rc = stats_static_ptr(&scratch)) return rc;
    return gqp_mala_stats(grad, energy, n_obj, batch_each, D, clip_grad, grad2_mean, z, scratch, stream);
}

extern "C" int gqp_mala_propose(const GqpMalaParams *p, const float *pose, const float *grad,
                                const float *grad2_mean, float *ema, int32_t *step, const int32_t *idx, int Bt, int D,
                                int n_c, int n_cand, const float *u_switch, const int32_t *idx_rand, uint64_t seed,
                                uint64_t iteration, float *pose_new, int32_t *idx_new, float *step_size_out,
                                void *stream) {
    if (!p || Bt < 0 || D <= 0 || n_c <= 0 || n_cand <= 0) return gqp_set_error(GQP_ERR_ARG, "gqp_mala_propose: bad argument");
    if (Bt == 0) return GQP_OK;
    if (!pose || !grad || !grad2_mean || !ema || !step || !idx || !pose_new || !idx_new)
        return gqp_set_error(GQP_ERR_ARG, "gqp_mala_propose: null pointer");
    if ((u_switch == nullptr) != (idx_rand == nullptr))
        return gqp_set_error(GQP_ERR_ARG, "gqp_mala_propose: u_switch and idx_rand must be injected together");
    GQP_LAUNCH(mala_propose_kernel, (Bt + 127) / 128, 128, 0, (cudaStream_t)stream, *p, pose, grad, grad2_mean, ema,
               step, idx, Bt, D, n_c, n_cand, u_switch, idx_rand, seed, iteration, pose_new, idx_new, step_size_out);
    GQP_CHECK_LAUNCH("gqp_mala_propose");
    return GQP_OK;
}

extern "C" int gqp_mala_accept(const GqpMalaParams *p, int Bt, int D, int n_c, int n_terms, const float *weights,
                               const float *losses_new, const float *z, const uint8_t *reset_mask,
                               const float *u_accept, uint64_t seed, uint64_t iteration, const int32_t *step,
                               const float *pose_new, const int32_t *idx_new, const float *grad_new, float *pose,
                               int32_t *idx, float *grad, float *energy, float *losses, float *new_energy_out,
                               uint8_t *accept_out, float *temperature_out, int after_reset, void *stream) {
    if (!p || Bt < 0 || D <= 0 || n_c <= 0 || n_terms <= 0) return gqp_set_error(GQP_ERR_ARG, "gqp_mala_accept: bad argument");
    if (Bt == 0) return GQP_OK;
    if (!weights || !losses_new || !step || !pose_new || !idx_new || !grad_new || !pose || !idx || !grad || !energy ||
        !losses)
        return gqp_set_error(GQP_ERR_ARG, "gqp_mala_accept: null pointer");
    GQP_LAUNCH(mala_accept_kernel, (Bt + 127) / 128, 128, 0, (cudaStream_t)stream, *p, Bt, D, n_c, n_terms, weights,
               losses_new, z, reset_mask, u_accept, seed, iteration, step, 0, pose_new, idx_new, grad_new, pose, idx,
               grad, energy, losses, new_energy_out, accept_out, temperature_out, after_reset != 0);
    GQP_CHECK_LAUNCH("gqp_mala_accept");
    return GQP_OK;
}

// ---- (f3) AnnealingDexGraspNet (optimizer.py:11-149; --optimizer dexgraspnet of scripts/fit.py:336-337) --------------------
// One RMS accumulator (D) and one step counter for the whole batch; no gradient clipping, no NaN handling, no per-row
// temperature: try_step :70-108, accept_step :110-138, reset_envs is a no-op (:148-149).
__global__ void __launch_bounds__(128)
anneal_propose_kernel(GqpMalaParams P, const float *__restrict__ pose, const float *__restrict__ grad,
                      const float *__restrict__ grad2_mean, const float *__restrict__ ema_in, float *__restrict__ ema_out,
                      int step, const int32_t *__restrict__ idx, int Bt, int D, int n_c, int n_cand,
                      const float *__restrict__ u_switch, const int32_t *__restrict__ idx_rand, uint64_t seed,
                      uint64_t iteration, float *__restrict__ pose_new, int32_t *__restrict__ idx_new,
                      float *__restrict__ step_size_out) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= Bt) return;
    const float s = P.step_size * powf(P.temperature_decay, (float)(step / P.stepsize_period));  // :81
    for (int d = 0; d < D; ++d) {
        const size_t i = (size_t)b * D + d;
        const float e = P.mu * grad2_mean[d] + (1.0f - P.mu) * ema_in[d];  // :84-86
        if (b == 0) ema_out[d] = e;  // ema_in stays intact for the other rows: the caller swaps the two buffers
        pose_new[i] = pose[i] - s * grad[i] / (sqrtf(e) + 1e-6f);  // :88-90
    }
    for (int c = 0; c < n_c; ++c) {  // :91-96
        const size_t i = (size_t)b * n_c + c;
        float u;
        int32_t r;
        if (u_switch) {
            u = u_switch[i];
            r = idx_rand[i];
        } else {
            Philox p = philox4x32((uint32_t)b, (uint32_t)c, (uint32_t)iteration, (uint32_t)(iteration >> 32),
                                  (uint32_t)seed, (uint32_t)(seed >> 32));
            u = u01(p.c[0]);
            r = (int32_t)(((uint64_t)p.c[1] * (uint64_t)n_cand) >> 32);
        }
        idx_new[i] = (u < P.switch_possibility) ? r : idx[i];
    }
    if (step_size_out && b == 0) step_size_out[0] = s;
}

extern "C" int gqp_anneal_propose(const GqpMalaParams *p, const float *pose, const float *grad, const float *grad2_mean,
                                  const float *ema_in, float *ema_out, int step, const int32_t *idx, int Bt, int D, int n_c,
                                  int n_cand, const float *u_switch, const int32_t *idx_rand, uint64_t seed,
                                  uint64_t iteration, float *pose_new, int32_t *idx_new, float *step_size_out,
                                  void *stream) {
    if (!p || Bt < 0 || D <= 0 || n_c <= 0 || n_cand <= 0 || step < 0)
        return gqp_set_error(GQP_ERR_ARG, "gqp_anneal_propose: bad argument");
    if (Bt == 0) return GQP_OK;
    if (!pose || !grad || !grad2_mean || !ema_in || !ema_out || !idx || !pose_new || !idx_new)
        return gqp_set_error(GQP_ERR_ARG, "gqp_anneal_propose: null pointer");
    if (ema_in == ema_out)
        return gqp_set_error(GQP_ERR_ARG, "gqp_anneal_propose: ema_in and ema_out must be different buffers");
    if ((u_switch == nullptr) != (idx_rand == nullptr))
        return gqp_set_error(GQP_ERR_ARG, "gqp_anneal_propose: u_switch and idx_rand must be injected together");
    GQP_LAUNCH(anneal_propose_kernel, (Bt + 127) / 128, 128, 0, (cudaStream_t)stream, *p, pose, grad, grad2_mean, ema_in,
               ema_out, step, idx, Bt, D, n_c, n_cand, u_switch, idx_rand, seed, iteration, pose_new, idx_new,
               step_size_out);
    GQP_CHECK_LAUNCH("gqp_anneal_propose");
    return GQP_OK;
}

extern "C" int gqp_anneal_accept(const GqpMalaParams *p, int Bt, int D, int n_c, int n_terms, const float *weights,
                                 const float *losses_new, const float *u_accept, uint64_t seed, uint64_t iteration,
                                 int step, const float *pose_new, const int32_t *idx_new, const float *grad_new,
                                 float *pose, int32_t *idx, float *grad, float *energy, float *losses,
                                 float *new_energy_out, uint8_t *accept_out, float *temperature_out, int after_reset,
                                 void *stream) {
    if (!p || Bt < 0 || D <= 0 || n_c <= 0 || n_terms <= 0 || step < 0)
        return gqp_set_error(GQP_ERR_ARG, "gqp_anneal_accept: bad argument");
    if (Bt == 0) return GQP_OK;
    if (!weights || !losses_new || !pose_new || !idx_new || !grad_new || !pose || !idx || !grad || !energy || !losses)
        return gqp_set_error(GQP_ERR_ARG, "gqp_anneal_accept: null pointer");
    GQP_LAUNCH(mala_accept_kernel, (Bt + 127) / 128, 128, 0, (cudaStream_t)stream, *p, Bt, D, n_c, n_terms, weights,
               losses_new, (const float *)nullptr, (const uint8_t *)nullptr, u_accept, seed, iteration,
               (const int32_t *)nullptr, step, pose_new, idx_new, grad_new, pose, idx, grad, energy, losses,
               new_energy_out, accept_out, temperature_out, after_reset != 0);
    GQP_CHECK_LAUNCH("gqp_anneal_accept");
    return GQP_OK;
}


// ---- (f4) entropy statistics of a batch (HandModel.joint_entropy / pose_entropy, hand_model.py:698-760; --log_entropy) ----
// 32-bin histograms (torch.histc: values outside [lo, hi] are dropped, hi falls into the last bin) over the rows of the batch:
// every joint between its limits, the translation components in [-0.1, 0.1], the rotation as the spherical coordinates
// (r, theta, phi) of its rotation vector.  A block per histogram, integer counts through shared-memory atomics (exact, any
// order), entropy -sum p log p by thread 0.  roma.rotmat_to_rotvec (absent here) is restated: unit quaternion with w >= 0,
// angle = 2 atan2(|v|, w), rotvec = angle * v / |v|.
#define ENT_BINS 32
__device__ __forceinline__ V3 ent_rotvec(V3 x, V3 y) {
    // utils/transforms.py:5-13 (special Gram-Schmidt), columns (x^, y^, z)
    const float nx = sqrtf(dot3(x, x));
    const V3 xh = x * (1.0f / nx);
    V3 yp = y - xh * dot3(xh, y);
    const V3 yh = yp * (1.0f / sqrtf(dot3(yp, yp)));
    const V3 zh = cross3(xh, yh);
    const float m00 = xh.x, m10 = xh.y, m20 = xh.z, m01 = yh.x, m11 = yh.y, m21 = yh.z, m02 = zh.x, m12 = zh.y, m22 = zh.z;
    float qx, qy, qz, qw;
    const float tr = m00 + m11 + m22;
    if (tr > 0.f) {
        const float s4 = 2.0f * sqrtf(tr + 1.0f);
        qw = 0.25f * s4; qx = (m21 - m12) / s4; qy = (m02 - m20) / s4; qz = (m10 - m01) / s4;
    } else if (m00 > m11 && m00 > m22) {
        const float s4 = 2.0f * sqrtf(1.0f + m00 - m11 - m22);
        qw = (m21 - m12) / s4; qx = 0.25f * s4; qy = (m01 + m10) / s4; qz = (m02 + m20) / s4;
    } else if (m11 > m22) {
        const float s4 = 2.0f * sqrtf(1.0f + m11 - m00 - m22);
        qw = (m02 - m20) / s4; qx = (m01 + m10) / s4; qy = 0.25f * s4; qz = (m12 + m21) / s4;
    } else {
        const float s4 = 2.0f * sqrtf(1.0f + m22 - m00 - m11);
        qw = (m10 - m01) / s4; qx = (m02 + m20) / s4; qy = (m12 + m21) / s4; qz = 0.25f * s4;
    }
    if (qw < 0.f) { qx = -qx; qy = -qy; qz = -qz; qw = -qw; }
    const float vn = sqrtf(qx * qx + qy * qy + qz * qz);
    const float angle = 2.0f * atan2f(vn, qw);
    const float half = 0.5f * angle;
    const float scale = fabsf(angle) <= 1e-3f ? 2.0f + angle * angle / 12.0f + 7.0f * angle * angle * angle * angle / 2880.0f
                                              : angle / sinf(half);
    return v3(scale * qx, scale * qy, scale * qz);
}

__global__ void __launch_bounds__(256)
entropy_hist_kernel(const float *__restrict__ pose, int Bt, int D, int n_dofs, const float *__restrict__ lower,
                    const float *__restrict__ upper, float *__restrict__ ent /* (n_dofs + 6) */) {
    __shared__ int hist[ENT_BINS];
    const int qn = blockIdx.x;  // [0, n_dofs): joints; then translation x, y, z; then r, theta, phi
    if (threadIdx.x < ENT_BINS) hist[threadIdx.x] = 0;
    __syncthreads();
    float lo, hi;
    if (qn < n_dofs) { lo = lower[qn]; hi = upper[qn]; }
    else if (qn < n_dofs + 3) { lo = -0.1f; hi = 0.1f; }
    else if (qn < n_dofs + 5) { lo = 0.f; hi = 3.14159265358979323846f; }
    else { lo = -3.14159265358979323846f; hi = 3.14159265358979323846f; }
    for (int b = threadIdx.x; b < Bt; b += blockDim.x) {
        const float *pr = pose + (size_t)b * D;
        float v;
        if (qn < n_dofs) v = pr[9 + qn];
        else if (qn < n_dofs + 3) v = pr[qn - n_dofs];
        else {
            const V3 rv = ent_rotvec(ld3(pr + 3), ld3(pr + 6));
            const float r = sqrtf(dot3(rv, rv));
            if (qn == n_dofs + 3) v = r;
            else if (qn == n_dofs + 4) v = acosf(rv.z / r);
            else v = (rv.y > 0.f ? 1.f : (rv.y < 0.f ? -1.f : 0.f)) * acosf(rv.x / sqrtf(rv.x * rv.x + rv.y * rv.y));
        }
        if (v >= lo && v <= hi) {  // NaN fails both: dropped, as histc does
            int bin = (int)((v - lo) * (float)ENT_BINS / (hi - lo));
            if (bin >= ENT_BINS) bin = ENT_BINS - 1;
            atomicAdd(&hist[bin], 1);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int total = 0;
        for (int k = 0; k < ENT_BINS; ++k) total += hist[k];
        float e = 0.f;
        for (int k = 0; k < ENT_BINS; ++k)
            if (hist[k] > 0) {
                const float pk = (float)hist[k] / (float)total;
                e -= pk * logf(pk);
            }
        ent[qn] = e;
    }
}

__global__ void entropy_final_kernel(const float *__restrict__ ent, int n_dofs, float *__restrict__ out3) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    float j = 0.f, t = 0.f, r = 0.f;
    for (int k = 0; k < n_dofs; ++k) j += ent[k] / (float)n_dofs;
    for (int k = 0; k < 3; ++k) {
        t += ent[n_dofs + k] / 3.0f;
        r += ent[n_dofs + 3 + k] / 3.0f;
    }
    out3[0] = j; out3[1] = t; out3[2] = r;
}

extern "C" int gqp_entropy_stats(const float *pose, int Bt, int n_dofs, const float *joints_lower, const float *joints_upper,
                                 float *scratch, float *out3, void *stream) {
    if (Bt <= 0 || n_dofs <= 0 || !pose || !joints_lower || !joints_upper || !scratch || !out3)
        return gqp_set_error(GQP_ERR_ARG, "gqp_entropy_stats: bad argument");
    GQP_LAUNCH(entropy_hist_kernel, n_dofs + 6, 256, 0, (cudaStream_t)stream, pose, Bt, 9 + n_dofs, n_dofs, joints_lower,
               joints_upper, scratch);
    GQP_CHECK_LAUNCH("gqp_entropy_stats");
    GQP_LAUNCH(entropy_final_kernel, 1, 32, 0, (cudaStream_t)stream, scratch, n_dofs, out3);
    GQP_CHECK_LAUNCH("gqp_entropy_stats/final");
    return GQP_OK;
}
