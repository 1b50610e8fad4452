// Per-element update rules shared by the local optimizer kernels (optim.cu) and the
// fused exchange kernels (comm.cu).  Every function is a faithful fp32 restatement of
// one rule of the reference; the citations give the Lua source each one mirrors.
#pragma once
#include "common.cuh"

namespace mpit {

enum RuleId : int {
  RULE_ADD = 0,       // p += g                      (asyncsgd/pserver.lua:91, BiCNN/pserver.lua:196-198)
  RULE_RMSPROP = 1,   // BiCNN/pserver.lua:123-136, BiCNN/optim-rmsprop-single.lua:24-32
  RULE_ADAM = 2,      // BiCNN/pserver.lua:140-155, BiCNN/optim-adam-single.lua:24-31
  RULE_ADAMAX = 3,    // BiCNN/pserver.lua:156-171, BiCNN/optim-adamax-single.lua:23-31
  RULE_ADAGRAD = 4,   // BiCNN/pserver.lua:172-183, BiCNN/optim-adagrad-single.lua:24-26
  RULE_ADADELTA = 5,  // BiCNN/pserver.lua:184-195, BiCNN/optim-adadelta-single.lua:23-27
};

// Scalars are prepared on the host in double precision and narrowed once, exactly as
// Lua narrows a number when it is handed to a FloatTensor method.
struct RuleHyper {
  float a;   // rmsprop: decay    adam: beta1      adamax: beta1   adagrad: -        adadelta: rho
  float b;   // rmsprop: 1-decay  adam: 1-beta1    adamax: 1-beta1 adagrad: -        adadelta: 1-rho
  float c;   // rmsprop: momentum adam: beta2      adamax: beta2   adagrad: -        adadelta: -
  float d;   // rmsprop: -        adam: 1-beta2    adamax: -       adagrad: -        adadelta: -
  float eps;
  float step;  // rmsprop: -lr    adam: -lr_t      adamax: -lr_t   adagrad: -clr     adadelta: -lr
};

// One element of a server/single rule.  s0..s3 are the rule's state vectors.
template <int RULE>
__device__ __forceinline__ void rule_apply(float& p, float g, float& s0, float& s1, float& s2,
                                           const RuleHyper& h) {
  if (RULE == RULE_ADD) {
    p = add_rn(p, g);
  } else if (RULE == RULE_RMSPROP) {
    // s0 = gradAccum, s1 = gradSqAccum, s2 = update
    s0 = add_rn(mul_rn(s0, h.a), mul_rn(h.b, g));
    s1 = add_rn(mul_rn(s1, h.a), mul_rn(h.b, mul_rn(g, g)));
    float rms = __fsqrt_rn(add_rn(sub_rn(s1, mul_rn(s0, s0)), h.eps));
    s2 = add_rn(mul_rn(s2, h.c), mul_rn(h.step, __fdiv_rn(g, rms)));
    p = add_rn(p, s2);
  } else if (RULE == RULE_ADAM) {
    // s0 = m, s1 = v
    s0 = add_rn(mul_rn(s0, h.a), mul_rn(h.b, g));
    s1 = add_rn(mul_rn(s1, h.c), mul_rn(h.d, mul_rn(g, g)));
    float den = add_rn(__fsqrt_rn(s1), h.eps);
    p = add_rn(p, mul_rn(h.step, __fdiv_rn(s0, den)));
  } else if (RULE == RULE_ADAMAX) {
    // s0 = m, s1 = u
    s0 = add_rn(mul_rn(s0, h.a), mul_rn(h.b, g));
    s1 = fmaxf(mul_rn(s1, h.c), add_rn(fabsf(g), h.eps));
    p = add_rn(p, mul_rn(h.step, __fdiv_rn(s0, s1)));
  } else if (RULE == RULE_ADAGRAD) {
    // s0 = paramVariance
    s0 = add_rn(s0, mul_rn(g, g));
    float std_ = add_rn(__fsqrt_rn(s0), h.eps);
    p = add_rn(p, mul_rn(h.step, __fdiv_rn(g, std_)));
  } else if (RULE == RULE_ADADELTA) {
    // s0 = paramVariance, s1 = accDelta
    s0 = add_rn(mul_rn(s0, h.a), mul_rn(h.b, mul_rn(g, g)));
    float std_ = __fsqrt_rn(add_rn(s0, h.eps));
    float delta = mul_rn(__fdiv_rn(__fsqrt_rn(add_rn(s1, h.eps)), std_), g);
    p = add_rn(p, mul_rn(h.step, delta));
    s1 = add_rn(mul_rn(s1, h.a), mul_rn(h.b, mul_rn(delta, delta)));
  }
}

// Refresh the tensor-core operand copies of one parameter: `nmaps` index tables of n entries each
// (forward layout, data-gradient layout, ...); -1 means "this parameter has no slot in that copy".
__device__ __forceinline__ void pack_store(float* wpk, const int* wmap, int nmaps, long long n, long long i,
                                           float x) {
  if (!wmap) { wpk[i] = x; return; }
  for (int m = 0; m < nmaps; ++m) {
    int o = wmap[(long long)m * n + i];
    if (o >= 0) wpk[o] = x;
  }
}

// Gradient post-processing of BiCNN/bicnn.lua:398-409 (L1, L2, clamp), applied while
// the gradient is in registers on its way into a push or a local rule.
__device__ __forceinline__ float grad_reg_clip(float g, float p, float l1, float l2, float clip) {
  if (l1 != 0.f) {
    float s = (p > 0.f) ? 1.f : ((p < 0.f) ? -1.f : 0.f);
    g = add_rn(g, mul_rn(s, l1));
  }
  if (l2 != 0.f) g = add_rn(g, mul_rn(p, l2));
  if (clip > 0.f) g = fminf(fmaxf(g, -clip), clip);
  return g;
}

// Nesterov/EAMSGD local update, post-backward half (asyncsgd/optim-msgd.lua:31-37,
// asyncsgd/optim-eamsgd.lua:37-44 and :70): x -= clr*g ; v -= clr*g ; x -= d.
__device__ __forceinline__ void msgd_apply_elem(float& x, float& v, float g, float d, float clr,
                                                float l2wd, bool use_mom, bool has_d) {
  if (l2wd != 0.f) g = add_rn(g, mul_rn(l2wd, x));
  float t = mul_rn(-clr, g);
  x = add_rn(x, t);
  if (use_mom) v = add_rn(v, t);
  if (has_d) x = sub_rn(x, d);
}

// Pre-forward half (asyncsgd/optim-msgd.lua:28-29): v *= mom ; x += v.
__device__ __forceinline__ void nesterov_lookahead_elem(float& x, float& v, float mom) {
  v = mul_rn(v, mom);
  x = add_rn(x, v);
}

}  // namespace mpit
