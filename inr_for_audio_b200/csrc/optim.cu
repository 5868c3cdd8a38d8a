// optim.cu -- torch.optim.Adam + ReduceLROnPlateau on the device (run.py:116-117,186-189).
//
// The reference syncs the host 3-4 times per step (loss.item() x3, scheduler.step(loss)); here the whole
// optimiser state lives in an inr_opt_state in device memory, so a fit step never touches the host:
//   k_adam      one pass over the flat params: 28 B/param (r g,p,m,v; w p,m,v) + 4 B to re-zero the gradient; the LAST block to
//               finish (atomic ticket) then runs the scheduler: loss = acc*scale, plateau state machine, t += 1, next bias
//               corrections in float64 -- every block has read the step's scalars by then, so one launch does both
#include "common.cuh"

namespace inr {

__device__ void opt_post(inr_opt_state* st, float loss_scale, float* loss_hist);

__global__ void __launch_bounds__(256) k_adam(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m,
                                              float* __restrict__ v, int64_t P, inr_opt_state* st, float loss_scale,
                                              float* loss_hist) {
  const float b1 = st->beta1, b2 = st->beta2, eps = st->adam_eps;
  const float step_size = st->step_size, bc2_sqrt = st->bc2_sqrt;
  const float w1 = st->omb1, w2 = st->omb2;   // (float)(1.0 - beta) formed in float64, as torch's Python scalars are
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * 4;
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < P; i += stride) {
    if (i + 3 < P && ((reinterpret_cast<uintptr_t>(p + i) & 15) == 0)) {
      float4 pp = *reinterpret_cast<float4*>(p + i), gg = *reinterpret_cast<float4*>(g + i);
      float4 mm = *reinterpret_cast<float4*>(m + i), vv = *reinterpret_cast<float4*>(v + i);
      float* pa = &pp.x; float* ga = &gg.x; float* ma = &mm.x; float* va = &vv.x;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        ma[q] = __fmaf_rn(w1, ga[q] - ma[q], ma[q]);            // exp_avg.lerp_(grad, 1-beta1)
        va[q] = va[q] * b2 + (w2 * ga[q]) * ga[q];              // exp_avg_sq.mul_(b2).addcmul_(g, g, 1-b2)
        const float denom = sqrtf(va[q]) / bc2_sqrt + eps;      // (sqrt(v)/sqrt(bc2)).add_(eps)
        pa[q] = pa[q] + (-step_size * ma[q]) / denom;           // addcdiv_(m, denom, value=-step_size)
      }
      *reinterpret_cast<float4*>(p + i) = pp;
      *reinterpret_cast<float4*>(m + i) = mm;
      *reinterpret_cast<float4*>(v + i) = vv;
      *reinterpret_cast<float4*>(g + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      for (int64_t j = i; j < min(P, i + 4); ++j) {
        const float gq = g[j];
        float mq = m[j], vq = v[j];
        mq = __fmaf_rn(w1, gq - mq, mq);
        vq = vq * b2 + (w2 * gq) * gq;
        const float denom = sqrtf(vq) / bc2_sqrt + eps;
        const float pq = p[j] + (-step_size * mq) / denom;
        p[j] = pq; m[j] = mq; v[j] = vq; g[j] = 0.f;
      }
    }
  }
  // ---- scheduler + next step's scalars, by the last block out
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned ticket = atomicAdd(reinterpret_cast<unsigned*>(&st->reserved0), 1u);
    if (ticket == gridDim.x - 1) {
      st->reserved0 = 0;
      opt_post(st, loss_scale, loss_hist);
    }
  }
}

__device__ void opt_prepare_next(inr_opt_state* st) {
  const int t = st->step + 1;
  const double bc1 = 1.0 - pow((double)st->beta1, (double)t);
  const double bc2 = 1.0 - pow((double)st->beta2, (double)t);
  st->step_size = (float)((double)st->lr / bc1);
  st->bc2_sqrt = (float)sqrt(bc2);
}

__global__ void k_opt_init(inr_opt_state* st, float lr, float min_lr, int hist_cap) {
  st->lr = lr; st->min_lr = min_lr; st->factor = 0.8f; st->threshold = 1e-4f; st->sched_eps = 1e-8f;
  st->beta1 = 0.9f; st->beta2 = 0.999f; st->adam_eps = 1e-8f;
  st->patience = 200; st->num_bad = 0; st->step = 0; st->hist_cap = hist_cap;
  st->best = __int_as_float(0x7f800000); st->last_loss = 0.f;
  st->gmax_acc = 0.f; st->gmax_prev = 0.f; st->hist_base = 0; st->reserved0 = 0;
  st->omb1 = (float)(1.0 - 0.9); st->omb2 = (float)(1.0 - 0.999);
  for (int i = 0; i < 4; ++i) st->loss_acc[i] = 0.0;
  opt_prepare_next(st);
}

// torch/optim/lr_scheduler.py ReduceLROnPlateau.step (mode='min', threshold_mode='rel', cooldown=0)
__device__ void opt_post(inr_opt_state* st, float loss_scale, float* loss_hist) {
  const float loss = (float)(st->loss_acc[0] * (double)loss_scale + st->loss_acc[1]);
  st->last_loss = loss;
  if (loss_hist && st->hist_cap > 0) {
    // non-negative ring index even when a caller rewound `step` below hist_base (teacher-forced tests, resume)
    int idx = (st->step - st->hist_base) % st->hist_cap;
    if (idx < 0) idx += st->hist_cap;
    loss_hist[idx] = loss;
  }
  st->gmax_prev = st->gmax_acc;
  st->gmax_acc = 0.f;
  // python: current < best * (1 - threshold), evaluated in float64 on a float32-valued loss
  const double thr = (double)st->best * (1.0 - (double)st->threshold);
  if ((double)loss < thr) { st->best = loss; st->num_bad = 0; } else { st->num_bad += 1; }
  if (st->num_bad > st->patience) {
    const double old_lr = (double)st->lr;
    const double new_lr = fmax(old_lr * (double)st->factor, (double)st->min_lr);
    if (old_lr - new_lr > (double)st->sched_eps) st->lr = (float)new_lr;
    st->num_bad = 0;
  }
  st->step += 1;
  for (int i = 0; i < 4; ++i) st->loss_acc[i] = 0.0;
  opt_prepare_next(st);
}

// One-collective exchange of a sharded step: the two float64 loss accumulators as fp32 hi + lo pairs behind the gradient
// block (engine.allreduce_fused), and back after the sum-allreduce (the spectral term, evaluated identically on every
// rank, is scaled by 1 / world).
__global__ void k_loss_to_f32(const inr_opt_state* st, float* tail) {
  for (int i = 0; i < 2; ++i) {
    const float hi = (float)st->loss_acc[i];
    tail[i] = hi;
    tail[2 + i] = (float)(st->loss_acc[i] - (double)hi);
  }
}
__global__ void k_loss_from_f32(inr_opt_state* st, float* tail, float spectral_scale) {
  st->loss_acc[0] = (double)tail[0] + (double)tail[2];
  st->loss_acc[1] = ((double)tail[1] + (double)tail[3]) * (double)spectral_scale;
  for (int i = 0; i < 4; ++i) tail[i] = 0.f;
}

int opt_init(inr_opt_state* st, float lr, float min_lr, int hist_cap, cudaStream_t s) {
  { LaunchScope ls("k_opt_init", s);
  k_opt_init<<<1, 1, 0, s>>>(st, lr, min_lr, hist_cap); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

int adam_step(float* params, float* grads, float* m, float* v, int64_t P, inr_opt_state* opt, float loss_scale,
              float* loss_hist, cudaStream_t s) {
  int blocks = (int)std::min<int64_t>(cdiv(P, 256 * 4), (int64_t)num_sms() * 8);
  if (blocks < 1) blocks = 1;
  { LaunchScope ls("k_adam", s);
  k_adam<<<blocks, 256, 0, s>>>(params, grads, m, v, P, opt, loss_scale, loss_hist); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}

}  // namespace inr

extern "C" int inr_opt_loss_to_f32(const inr_opt_state* opt, float* tail4, void* stream) {
  using namespace inr;
  INR_REQUIRE(opt && tail4, INR_E_BADARG, "inr_opt_loss_to_f32: null argument");
  { LaunchScope ls("k_loss_xchg", (cudaStream_t)stream);
    k_loss_to_f32<<<1, 1, 0, (cudaStream_t)stream>>>(opt, tail4); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}
extern "C" int inr_opt_loss_from_f32(inr_opt_state* opt, float* tail4, float spectral_scale, void* stream) {
  using namespace inr;
  INR_REQUIRE(opt && tail4, INR_E_BADARG, "inr_opt_loss_from_f32: null argument");
  { LaunchScope ls("k_loss_xchg", (cudaStream_t)stream);
    k_loss_from_f32<<<1, 1, 0, (cudaStream_t)stream>>>(opt, tail4, spectral_scale); }
  INR_LAUNCH_CHECK();
  return INR_OK;
}
