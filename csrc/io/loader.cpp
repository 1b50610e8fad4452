// Background mini-batch assembler: a worker thread gathers shuffled samples of an in-memory dataset
// into a ring of caller-provided (pinned) batch slots, together with the labels and the per-sample
// random-crop offsets, so that the training thread only hands finished slots to the GPU.
//
// What it stands in for: the reference assembles every mini-batch inside the training loop, one
// sample at a time, in Lua -- per-epoch `torch.randperm` (/root/reference/asyncsgd/goot.lua:164),
// then for each of the 128 samples a random 28x28 crop position and an `image.crop`
// (goot.lua:172-184) -- on the same thread that runs the optimizer.  Here the crop itself is free
// (it is a TMA box offset in the first convolution), so a batch is: B sample copies,
// B labels, 2B crop offsets; at 0.13 ms per training step doing that on the Python thread would
// cost as much as the step itself.
//
// Threading contract: one producer (the worker) and one consumer (the thread that calls next /
// release).  Slots cycle free -> filling -> ready -> handed out -> free.  The generator is a
// SplitMix64 stream owned by the worker: batches are reproducible for a given seed, independent of
// timing.
#include "mpit_msg.h"

#include <stdint.h>
#include <string.h>

#include <condition_variable>
#include <deque>
#include <mutex>
#include <numeric>
#include <thread>
#include <vector>

namespace {

struct SplitMix64 {
  uint64_t s;
  uint64_t next() {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  // unbiased enough for shuffling / crop positions: 64-bit multiply-shift
  uint64_t below(uint64_t n) { return (uint64_t)(((unsigned __int128)next() * n) >> 64); }
};

struct Loader {
  const uint8_t* images;
  const int32_t* labels;
  long long n;
  long long sample_bytes;
  int batch, nslots, crop_range, shuffle;
  uint8_t* ring_img;
  int32_t* ring_lab;
  int32_t* ring_crop;
  SplitMix64 rng;
  std::vector<long long> perm;
  long long cursor = 0;
  long long epoch = 0;
  std::vector<long long> slot_epoch;
  std::mutex mu;
  std::condition_variable cv;
  std::deque<int> ready, free_;
  bool stop = false;
  std::thread th;

  void reshuffle() {
    if (!shuffle) return;
    for (long long i = n - 1; i > 0; --i) std::swap(perm[(size_t)i], perm[(size_t)rng.below((uint64_t)i + 1)]);
  }

  void fill(int slot) {
    if (cursor + batch > n) {          // a trailing partial batch is dropped, as goot.lua's trsize loop does
      cursor = 0;
      ++epoch;
      reshuffle();
    }
    uint8_t* dst = ring_img + (size_t)slot * batch * sample_bytes;
    int32_t* lab = ring_lab + (size_t)slot * batch;
    int32_t* crop = ring_crop + (size_t)slot * batch * 2;
    for (int i = 0; i < batch; ++i) {
      const long long idx = perm[(size_t)(cursor + i)];
      memcpy(dst + (size_t)i * sample_bytes, images + (size_t)idx * sample_bytes, (size_t)sample_bytes);
      lab[i] = labels[idx];
      crop[2 * i] = crop_range > 1 ? (int32_t)rng.below((uint64_t)crop_range) : 0;       // y
      crop[2 * i + 1] = crop_range > 1 ? (int32_t)rng.below((uint64_t)crop_range) : 0;   // x
    }
    cursor += batch;
    slot_epoch[(size_t)slot] = epoch;
  }

  void run() {
    for (;;) {
      int slot;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return stop || !free_.empty(); });
        if (stop) return;
        slot = free_.front();
        free_.pop_front();
      }
      fill(slot);
      {
        std::lock_guard<std::mutex> lk(mu);
        ready.push_back(slot);
      }
      cv.notify_all();
    }
  }
};

}  // namespace

extern "C" {

int mpit_loader_create(const unsigned char* images, const int* labels, long long n_samples, long long sample_bytes,
                       int batch, int nslots, unsigned char* ring_images, int* ring_labels, int* ring_crop,
                       int crop_range, unsigned long long seed, int shuffle, void** handle) {
  if (!images || !labels || !ring_images || !ring_labels || !ring_crop || !handle) return MPIT_ERR_BUFFER;
  if (batch <= 0 || n_samples < batch || sample_bytes <= 0 || nslots < 2 || crop_range < 0) return MPIT_ERR_ARG;
  Loader* L = new Loader();
  L->images = images;
  L->labels = labels;
  L->n = n_samples;
  L->sample_bytes = sample_bytes;
  L->batch = batch;
  L->nslots = nslots;
  L->crop_range = crop_range;
  L->shuffle = shuffle;
  L->ring_img = ring_images;
  L->ring_lab = ring_labels;
  L->ring_crop = ring_crop;
  L->rng.s = seed;
  L->perm.resize((size_t)n_samples);
  std::iota(L->perm.begin(), L->perm.end(), 0LL);
  L->slot_epoch.assign((size_t)nslots, 0);
  L->reshuffle();
  for (int s = 0; s < nslots; ++s) L->free_.push_back(s);
  L->th = std::thread([L] { L->run(); });
  *handle = L;
  return MPIT_SUCCESS;
}

// Blocks until a batch is ready; *slot indexes the ring, *epoch counts completed passes over the data.
int mpit_loader_next(void* handle, int* slot, long long* epoch) {
  Loader* L = (Loader*)handle;
  if (!L) return MPIT_ERR_ARG;
  std::unique_lock<std::mutex> lk(L->mu);
  L->cv.wait(lk, [&] { return !L->ready.empty(); });
  *slot = L->ready.front();
  L->ready.pop_front();
  if (epoch) *epoch = L->slot_epoch[(size_t)*slot];
  return MPIT_SUCCESS;
}

// Hand a slot back to the worker once its contents have been consumed (copied to the device).
int mpit_loader_release(void* handle, int slot) {
  Loader* L = (Loader*)handle;
  if (!L || slot < 0 || slot >= L->nslots) return MPIT_ERR_ARG;
  {
    std::lock_guard<std::mutex> lk(L->mu);
    L->free_.push_back(slot);
  }
  L->cv.notify_all();
  return MPIT_SUCCESS;
}

int mpit_loader_destroy(void* handle) {
  Loader* L = (Loader*)handle;
  if (!L) return MPIT_ERR_ARG;
  {
    std::lock_guard<std::mutex> lk(L->mu);
    L->stop = true;
  }
  L->cv.notify_all();
  if (L->th.joinable()) L->th.join();
  delete L;
  return MPIT_SUCCESS;
}

}  // extern "C"
