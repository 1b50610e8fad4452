// Local optimizer kernels (sm_100a).  One fused elementwise pass per rule instead of the
// reference's chains of 4-8 TH BLAS-1 calls with temporaries (BiCNN/pserver.lua:123-195).
// All of these are HBM/L2-streaming kernels: float4 main path, scalar tail, no smem.
#include "mpit_b200.h"
#include "optim_math.cuh"

namespace mpit {

int g_pdl_enabled = 1;
int g_x3_mode = 0;

// ---------------------------------------------------------------------------------------
// local_step: [apply(t)] -> [lookahead(t+1)] -> [pack] in one pass.
// ---------------------------------------------------------------------------------------
struct LocalStepArgs {
  float* x;
  float* v;            // may be null when mom is never used
  const float* g;      // gradient (flat order, or packed when gmap != null)
  const int* gmap;     // flat index -> offset in g (null: identity)
  const float* d;      // elastic difference to subtract after the step (null: none)
  float* wpk;          // packed weights to refresh (null: none)
  const int* wmap;     // flat index -> offset in wpk; nmaps tables of n entries, -1 = skip
  int nmaps;
  long long n;
  float clr, l2wd, mom_next;
  int do_apply, use_mom, do_lookahead;
};

__global__ void __launch_bounds__(256) local_step_kernel(LocalStepArgs a) {
  pdl_trigger();
  pdl_wait();
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
    float x = a.x[i];
    float v = a.use_mom ? a.v[i] : 0.f;
    if (a.do_apply) {
      float g = a.gmap ? a.g[a.gmap[i]] : a.g[i];
      float d = a.d ? a.d[i] : 0.f;
      msgd_apply_elem(x, v, g, d, a.clr, a.l2wd, a.use_mom != 0, a.d != nullptr);
    }
    if (a.do_lookahead && a.use_mom) nesterov_lookahead_elem(x, v, a.mom_next);
    a.x[i] = x;
    if (a.use_mom) a.v[i] = v;
    if (a.wpk) pack_store(a.wpk, a.wmap, a.nmaps, a.n, i, x);
  }
}

// Vectorised variant for the map-free case (bandwidth sweeps, big flat tensors).
__global__ void __launch_bounds__(256) local_step_vec_kernel(LocalStepArgs a) {
  pdl_trigger();
  pdl_wait();
  long long n4 = a.n >> 2;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 x = ld4(a.x + 4 * i);
    float4 v = a.use_mom ? ld4(a.v + 4 * i) : make_float4(0, 0, 0, 0);
    if (a.do_apply) {
      float4 g = ld4_stream(a.g + 4 * i);
      float4 d = a.d ? ld4_stream(a.d + 4 * i) : make_float4(0, 0, 0, 0);
      bool um = a.use_mom != 0, hd = a.d != nullptr;
      msgd_apply_elem(x.x, v.x, g.x, d.x, a.clr, a.l2wd, um, hd);
      msgd_apply_elem(x.y, v.y, g.y, d.y, a.clr, a.l2wd, um, hd);
      msgd_apply_elem(x.z, v.z, g.z, d.z, a.clr, a.l2wd, um, hd);
      msgd_apply_elem(x.w, v.w, g.w, d.w, a.clr, a.l2wd, um, hd);
    }
    if (a.do_lookahead && a.use_mom) {
      nesterov_lookahead_elem(x.x, v.x, a.mom_next);
      nesterov_lookahead_elem(x.y, v.y, a.mom_next);
      nesterov_lookahead_elem(x.z, v.z, a.mom_next);
      nesterov_lookahead_elem(x.w, v.w, a.mom_next);
    }
    st4(a.x + 4 * i, x);
    if (a.use_mom) st4(a.v + 4 * i, v);
    if (a.wpk) st4(a.wpk + 4 * i, x);
  }
  // scalar tail
  long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < a.n) {
    float x = a.x[i];
    float v = a.use_mom ? a.v[i] : 0.f;
    if (a.do_apply) msgd_apply_elem(x, v, a.g[i], a.d ? a.d[i] : 0.f, a.clr, a.l2wd, a.use_mom != 0, a.d != nullptr);
    if (a.do_lookahead && a.use_mom) nesterov_lookahead_elem(x, v, a.mom_next);
    a.x[i] = x;
    if (a.use_mom) a.v[i] = v;
    if (a.wpk) a.wpk[i] = x;
  }
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---------------------------------------------------------------------------------------
// Server / single rules
// ---------------------------------------------------------------------------------------
struct RuleArgs {
  float* p;
  const float* g;
  const int* gmap;
  float* s0;
  float* s1;
  float* s2;
  float* wpk;
  const int* wmap;
  int nmaps;
  long long n;
  RuleHyper h;
  float l1, l2, clip;  // optional gradient prologue (bicnn.lua:398-409)
  float gscale;        // gradient pre-scale (e.g. 1/batch); 1 = none
};

template <int RULE>
__global__ void __launch_bounds__(256) rule_kernel(RuleArgs a) {
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
    float p = a.p[i];
    float g = a.gmap ? a.g[a.gmap[i]] : a.g[i];
    if (a.gscale != 1.f) g = mul_rn(g, a.gscale);
    g = grad_reg_clip(g, p, a.l1, a.l2, a.clip);
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
    if (RULE != RULE_ADD) s0 = a.s0[i];
    if (RULE == RULE_RMSPROP || RULE == RULE_ADAM || RULE == RULE_ADAMAX || RULE == RULE_ADADELTA) s1 = a.s1[i];
    if (RULE == RULE_RMSPROP) s2 = a.s2[i];
    rule_apply<RULE>(p, g, s0, s1, s2, a.h);
    a.p[i] = p;
    if (RULE != RULE_ADD) a.s0[i] = s0;
    if (RULE == RULE_RMSPROP || RULE == RULE_ADAM || RULE == RULE_ADAMAX || RULE == RULE_ADADELTA) a.s1[i] = s1;
    if (RULE == RULE_RMSPROP) a.s2[i] = s2;
    if (a.wpk) pack_store(a.wpk, a.wmap, a.nmaps, a.n, i, p);
  }
}

// ---------------------------------------------------------------------------------------
// Downpour client-side accumulate (asyncsgd/optim-downpour.lua:26-45):
//   g' = (g + l2wd*x) * (-lr) ; acc += g' ; [x += g' when not a sync step]
// ---------------------------------------------------------------------------------------
struct DownpourArgs {
  float* x;
  float* acc;
  const float* g;
  const int* gmap;
  float* wpk;
  const int* wmap;
  int nmaps;
  long long n;
  float neg_lr, l2wd;
  int move_local;
};

__global__ void __launch_bounds__(256) downpour_accum_kernel(DownpourArgs a) {
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
    float g = a.gmap ? a.g[a.gmap[i]] : a.g[i];
    float x = a.x[i];
    if (a.l2wd != 0.f) g = add_rn(g, mul_rn(a.l2wd, x));
    g = mul_rn(g, a.neg_lr);
    a.acc[i] = add_rn(a.acc[i], g);
    if (a.move_local) {
      x = add_rn(x, g);
      a.x[i] = x;
      if (a.wpk) pack_store(a.wpk, a.wmap, a.nmaps, a.n, i, x);
    }
  }
}

__global__ void __launch_bounds__(256) grad_reg_clip_kernel(float* g, const float* p, long long n, float l1,
                                                            float l2, float clip) {
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    g[i] = grad_reg_clip(g[i], p[i], l1, l2, clip);
}

__global__ void __launch_bounds__(256) gather_map_kernel(float* dst, const float* src, const int* map, long long n,
                                                         int scatter) {
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (scatter) dst[map[i]] = src[i];
    else dst[i] = src[map[i]];
  }
}

}  // namespace mpit

using namespace mpit;

extern "C" {

int mpit_set_pdl(int enabled) {
  int old = g_pdl_enabled;
  if (enabled >= 0) g_pdl_enabled = enabled ? 1 : 0;
  return old;
}

int mpit_set_tf32x3(int mode) {
  int old = g_x3_mode;
  if (mode >= 0 && mode <= 2) g_x3_mode = mode;
  return old;
}

int mpit_local_step(float* x, float* v, const float* g, const int* gmap, const float* d, float* wpk,
                    const int* wmap, int nmaps, long long n, float clr, float l2wd, float mom_next, int do_apply,
                    int use_mom, int do_lookahead, void* stream) {
  if (n <= 0) return 0;
  LocalStepArgs a{x, v, g, gmap, d, wpk, wmap, nmaps, n, clr, l2wd, mom_next, do_apply, use_mom, do_lookahead};
  cudaStream_t st = (cudaStream_t)stream;
  bool vec = !gmap && !wmap && aligned16(x) && (!v || aligned16(v)) && (!g || aligned16(g)) &&
             (!d || aligned16(d)) && (!wpk || aligned16(wpk));
  if (vec) {
    int grid = ew_grid(n, 256, 8);
    MPIT_CUDA_TRY(launch_k(local_step_vec_kernel, dim3(grid), dim3(256), 0, st, a));
  } else {
    int grid = ew_grid(n, 256, 4);
    MPIT_CUDA_TRY(launch_k(local_step_kernel, dim3(grid), dim3(256), 0, st, a));
  }
  return 0;
}

int mpit_rule_step(int rule, float* p, const float* g, const int* gmap, float* s0, float* s1, float* s2,
                   float* wpk, const int* wmap, int nmaps, long long n, const float* hyper6, float l1,
                   float l2, float clip, float gscale, void* stream) {
  if (n <= 0) return 0;
  RuleArgs a{};
  a.p = p; a.g = g; a.gmap = gmap; a.s0 = s0; a.s1 = s1; a.s2 = s2; a.wpk = wpk; a.wmap = wmap; a.nmaps = nmaps; a.n = n;
  a.h = RuleHyper{hyper6[0], hyper6[1], hyper6[2], hyper6[3], hyper6[4], hyper6[5]};
  a.l1 = l1; a.l2 = l2; a.clip = clip; a.gscale = gscale;
  cudaStream_t st = (cudaStream_t)stream;
  int grid = ew_grid(n, 256, 4);
  switch (rule) {
    case RULE_ADD: rule_kernel<RULE_ADD><<<grid, 256, 0, st>>>(a); break;
    case RULE_RMSPROP: rule_kernel<RULE_RMSPROP><<<grid, 256, 0, st>>>(a); break;
    case RULE_ADAM: rule_kernel<RULE_ADAM><<<grid, 256, 0, st>>>(a); break;
    case RULE_ADAMAX: rule_kernel<RULE_ADAMAX><<<grid, 256, 0, st>>>(a); break;
    case RULE_ADAGRAD: rule_kernel<RULE_ADAGRAD><<<grid, 256, 0, st>>>(a); break;
    case RULE_ADADELTA: rule_kernel<RULE_ADADELTA><<<grid, 256, 0, st>>>(a); break;
    default: return (int)cudaErrorInvalidValue;
  }
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_downpour_accum(float* x, float* acc, const float* g, const int* gmap, float* wpk, const int* wmap,
                        int nmaps, long long n, float neg_lr, float l2wd, int move_local, void* stream) {
  if (n <= 0) return 0;
  DownpourArgs a{x, acc, g, gmap, wpk, wmap, nmaps, n, neg_lr, l2wd, move_local};
  downpour_accum_kernel<<<ew_grid(n, 256, 4), 256, 0, (cudaStream_t)stream>>>(a);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_grad_reg_clip(float* g, const float* p, long long n, float l1, float l2, float clip, void* stream) {
  if (n <= 0) return 0;
  grad_reg_clip_kernel<<<ew_grid(n, 256, 4), 256, 0, (cudaStream_t)stream>>>(g, p, n, l1, l2, clip);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_map_copy(float* dst, const float* src, const int* map, long long n, int scatter, void* stream) {
  if (n <= 0) return 0;
  gather_map_kernel<<<ew_grid(n, 256, 4), 256, 0, (cudaStream_t)stream>>>(dst, src, map, n, scatter);
  MPIT_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
