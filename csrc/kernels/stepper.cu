// Host side of one end-to-end training step, in native code: input staging on a copy stream into a
// double-buffered device area, the input-conversion kernel, the read-back of the loss with a lag, and
// the stream / event choreography between them.  The optimizer step itself (a CUDA-graph replay)
// is launched by the caller between stage() and finish().
//
// The reference's step loop lives in Lua and spends its time per sample (/root/reference/asyncsgd/
// goot.lua:172-225); with a 0.13 ms device step the analogous Python loop here (a dozen torch stream,
// event and copy calls per step) was the bottleneck of the end-to-end path, so those calls are issued
// from one C function instead.
//
// Pipeline (k = step index, b = k & 1, LAG = loss read-back lag):
//   copy stream : wait in_free[b] (step k-2 consumed the slot) -> H2D images, labels, crops -> record in_ready[b]
//   main stream : wait in_ready[b] -> prep kernel (uint8 -> network input, labels, crops; ALSO stores the loss of
//                 step k-1 into its host slot) -> record in_free[b], loss_evt[k-1] -> [caller: optimizer step k]
//   host        : returns after in_ready[b] (caller's pinned buffers reusable); the loss of step k-LAG is
//                 returned by stage() of step k.
//
// Loss read-back.  A 4-byte cudaMemcpyAsync per step on the compute stream cost 6-10 us of a 130 us step (measured:
// 0.141 ms/step with it, 0.130 without; the image H2D on the copy stream and the host-side wait were free).  The
// pinned loss slots are therefore written by the next step's prep kernel through their device-side alias (pinned
// memory is device-accessible under unified addressing); only the newest step, which has no successor yet, is
// fetched with a copy when somebody asks for it.  If the alias cannot be obtained the copy path is used throughout.
//
// No device memory is allocated here: staging buffers and the pinned loss slots belong to the caller.
#include <cuda_runtime.h>
#include <stdint.h>

#include "mpit_b200.h"

namespace {

constexpr int kMaxLag = 6;

struct Stepper {
  int lag;
  cudaStream_t copy;
  cudaEvent_t in_ready[2], in_free[2], loss_evt[kMaxLag + 1];
  unsigned char* d_img[2];
  int* d_lab[2];
  int* d_crop[2];
  float* loss_host;       // pinned, [(lag+1)][2]
  float* loss_host_dev;   // the same memory as the device sees it (nullptr: copy path)
  const float* loss_dev = nullptr;   // where the newest step left its loss
  cudaStream_t last_main = nullptr;
  bool pending = false;   // the newest step's loss has not been sent to the host yet
  long long img_bytes;
  int B, SH, SW, cin;
  long long k = 0, ret = 0;
};

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return (int)e_; } while (0)

// Loss of the newest finished step, fetched with a copy (nobody has launched a following step yet).
int flush_pending(Stepper* s) {
  if (!s->pending) return 0;
  const int slot = (int)((s->k - 1) % (s->lag + 1));
  CK(cudaMemcpyAsync(s->loss_host + 2 * slot, s->loss_dev, 4, cudaMemcpyDeviceToHost, s->last_main));
  CK(cudaEventRecord(s->loss_evt[slot], s->last_main));
  s->pending = false;
  return 0;
}

}  // namespace

extern "C" int mpit_stepper_create(void** handle, int lag, void* d_img0, void* d_img1, int* d_lab0, int* d_lab1,
                                   int* d_crop0, int* d_crop1, float* loss_host_pinned, long long img_bytes, int B,
                                   int SH, int SW, int cin) {
  if (!handle || lag < 0 || lag > kMaxLag || !d_img0 || !d_img1 || !loss_host_pinned || img_bytes <= 0) return MPIT_E_ARG;
  Stepper* s = new Stepper();
  s->lag = lag;
  s->d_img[0] = (unsigned char*)d_img0; s->d_img[1] = (unsigned char*)d_img1;
  s->d_lab[0] = d_lab0; s->d_lab[1] = d_lab1;
  s->d_crop[0] = d_crop0; s->d_crop[1] = d_crop1;
  s->loss_host = loss_host_pinned;
  void* alias = nullptr;
  if (cudaHostGetDevicePointer(&alias, loss_host_pinned, 0) == cudaSuccess) s->loss_host_dev = (float*)alias;
  else { s->loss_host_dev = nullptr; (void)cudaGetLastError(); }
  s->img_bytes = img_bytes;
  s->B = B; s->SH = SH; s->SW = SW; s->cin = cin;
  CK(cudaStreamCreateWithFlags(&s->copy, cudaStreamNonBlocking));
  for (int i = 0; i < 2; ++i) {
    CK(cudaEventCreateWithFlags(&s->in_ready[i], cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&s->in_free[i], cudaEventDisableTiming));
  }
  for (int i = 0; i <= lag; ++i) CK(cudaEventCreateWithFlags(&s->loss_evt[i], cudaEventDisableTiming));
  *handle = s;
  return 0;
}

// Stage the inputs of the next step.  Host pointers must be page-locked for the copies to be asynchronous.
// *have_prev = 1 and *prev_loss_sum = summed loss of step k-LAG when that step exists (waits for it).
extern "C" int mpit_stepper_stage(void* handle, const void* img, const int* labels, const int* crop, float* x0_dst,
                                  int* labels_dst, int* crop_dst, void* main_stream, float* prev_loss_sum,
                                  int* have_prev) {
  Stepper* s = (Stepper*)handle;
  if (!s || !img || !labels || !crop) return MPIT_E_ARG;
  cudaStream_t main = (cudaStream_t)main_stream;
  *have_prev = 0;
  if (s->k - s->ret >= s->lag) {
    const int slot = (int)(s->ret % (s->lag + 1));
    if (s->ret == s->k - 1) { int rc = flush_pending(s); if (rc) return rc; }    // LAG = 0 or 1: no successor yet
    CK(cudaEventSynchronize(s->loss_evt[slot]));
    *prev_loss_sum = s->loss_host[2 * slot];
    *have_prev = 1;
    s->ret++;
  }
  const int b = (int)(s->k & 1);
  if (s->k >= 2) CK(cudaStreamWaitEvent(s->copy, s->in_free[b], 0));
  CK(cudaMemcpyAsync(s->d_img[b], img, (size_t)s->img_bytes, cudaMemcpyHostToDevice, s->copy));
  CK(cudaMemcpyAsync(s->d_lab[b], labels, (size_t)s->B * 4, cudaMemcpyHostToDevice, s->copy));
  CK(cudaMemcpyAsync(s->d_crop[b], crop, (size_t)s->B * 8, cudaMemcpyHostToDevice, s->copy));
  CK(cudaEventRecord(s->in_ready[b], s->copy));
  CK(cudaStreamWaitEvent(main, s->in_ready[b], 0));
  const bool forward = s->pending && s->loss_host_dev != nullptr;
  const int prev_slot = (int)((s->k - 1 + (s->lag + 1)) % (s->lag + 1));
  int rc = mpit_prep_u8_fwd(s->d_img[b], x0_dst, s->B, s->SH, s->SW, s->cin, s->d_lab[b], labels_dst, s->d_crop[b],
                            crop_dst, forward ? s->loss_dev : nullptr,
                            forward ? s->loss_host_dev + 2 * prev_slot : nullptr, forward ? 1 : 0, main_stream);
  if (rc) return rc;
  CK(cudaEventRecord(s->in_free[b], main));
  if (forward) {
    CK(cudaEventRecord(s->loss_evt[prev_slot], main));
    s->pending = false;
  }
  return 0;
}

// After the optimizer step has been enqueued on `main_stream`: note where its loss will be and wait until the
// inputs staged by the matching stage() call have left the host buffers.
extern "C" int mpit_stepper_finish(void* handle, const float* loss_dev, void* main_stream) {
  Stepper* s = (Stepper*)handle;
  if (!s || !loss_dev) return MPIT_E_ARG;
  cudaStream_t main = (cudaStream_t)main_stream;
  const int b = (int)(s->k & 1);
  s->loss_dev = loss_dev;
  s->last_main = main;
  s->pending = true;
  s->k++;
  if (!s->loss_host_dev) { int rc = flush_pending(s); if (rc) return rc; }     // copy path
  CK(cudaEventSynchronize(s->in_ready[b]));
  return 0;
}

// Summed loss of step `index` (0-based, counting stage/finish pairs).  Only the last LAG+1 steps are kept.
// index < 0: the most recent step.  Does not change which loss stage() will hand back next.
extern "C" int mpit_stepper_loss(void* handle, long long index, float* loss_sum) {
  Stepper* s = (Stepper*)handle;
  if (!s || s->k == 0) return MPIT_E_ARG;
  if (index < 0) index = s->k - 1;
  if (index >= s->k || index < s->k - (s->lag + 1)) return MPIT_E_ARG;
  if (index == s->k - 1) { int rc = flush_pending(s); if (rc) return rc; }
  const int slot = (int)(index % (s->lag + 1));
  CK(cudaEventSynchronize(s->loss_evt[slot]));
  *loss_sum = s->loss_host[2 * slot];
  return 0;
}

// Counters: *issued = steps finished, *returned = steps whose loss stage() has handed back.
extern "C" int mpit_stepper_counters(void* handle, long long* issued, long long* returned, int mark_returned) {
  Stepper* s = (Stepper*)handle;
  if (!s) return MPIT_E_ARG;
  if (mark_returned) s->ret = s->k;
  *issued = s->k;
  *returned = s->ret;
  return 0;
}

extern "C" int mpit_stepper_destroy(void* handle) {
  Stepper* s = (Stepper*)handle;
  if (!s) return MPIT_E_ARG;
  cudaStreamSynchronize(s->copy);
  for (int i = 0; i < 2; ++i) { cudaEventDestroy(s->in_ready[i]); cudaEventDestroy(s->in_free[i]); }
  for (int i = 0; i <= s->lag; ++i) cudaEventDestroy(s->loss_evt[i]);
  cudaStreamDestroy(s->copy);
  delete s;
  return 0;
}
