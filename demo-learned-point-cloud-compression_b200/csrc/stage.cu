// (7) Host -> device staging of raw frames: pageable host arrays -> pinned memory -> device, in pieces, by a
// small pool of native threads.  Each worker copies a piece into the pinned buffer and enqueues that piece's
// DMA itself, so the bus is busy while the next pieces are still being copied.  The reference hands numpy
// arrays to torch.tensor(..., device=) (sender/encoder/codec_pipeline.py:252-266: pageable, synchronous);
// doing the piecewise copy from Python threads serialised on the interpreter lock (one acquisition per piece).
#include <atomic>
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

#include <unistd.h>

#include "common.cuh"

// host_convert.cpp (declared in include/pcc_b200.h)

namespace {

struct Piece {
  const uint8_t* src;
  uint8_t* pinned;
  uint8_t* dev;
  int64_t len;       // source bytes
  int64_t out_len;   // bytes written to pinned / device (len, or len / 4 when converting f64 -> bf16)
};

struct Batch {
  std::vector<Piece> pieces;
  std::atomic<int64_t> next{0};
  std::atomic<int> err{0};
  int device = 0;
  int convert = 0;
  cudaStream_t stream = nullptr;
  int pending = 0;          // workers still inside this batch (guarded by Pool::mu)
};

class Pool {
 public:
  // one pool per process: a child created by fork() inherits the object but not its threads, so it gets a
  // fresh pool (the inherited one is abandoned, never destroyed: its mutex state is unknown)
  static Pool& get() {
    static std::mutex guard;
    static Pool* pool = nullptr;
    static pid_t owner = 0;
    std::lock_guard<std::mutex> lk(guard);
    if (!pool || owner != getpid()) {
      pool = new Pool();
      owner = getpid();
    }
    return *pool;
  }
  // runs the batch on `threads` workers (the caller is one of them) and returns when every piece is enqueued
  int run(Batch& b, int threads) {
    const int helpers = threads - 1;
    {
      std::lock_guard<std::mutex> lk(mu);
      while (nworkers < helpers) {
        std::thread([this] { loop(); }).detach();
        ++nworkers;
      }
      b.pending = helpers;
      for (int i = 0; i < helpers; ++i) queue.push_back(&b);
    }
    cv.notify_all();
    work(b);
    std::unique_lock<std::mutex> lk(mu);
    done.wait(lk, [&] { return b.pending == 0; });
    return b.err.load();
  }

 private:
  Pool() = default;     // lives until the process exits; its idle workers are detached
  static void work(Batch& b) {
    int cur = -1;
    for (;;) {
      const int64_t i = b.next.fetch_add(1);
      if (i >= (int64_t)b.pieces.size()) break;
      const Piece& p = b.pieces[(size_t)i];
      if (b.convert)
        pcc_host_f64_to_bf16((const double*)p.src, p.len / 8, (uint16_t*)p.pinned);
      else
        pcc_host_copy_stream(p.src, p.len, p.pinned);
      if (cur != b.device) {
        if (cudaSetDevice(b.device) != cudaSuccess) b.err.store(1);
        cur = b.device;
      }
      if (cudaMemcpyAsync(p.dev, p.pinned, (size_t)p.out_len, cudaMemcpyHostToDevice, b.stream) != cudaSuccess) b.err.store(1);
    }
  }
  void loop() {
    for (;;) {
      Batch* b = nullptr;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return !queue.empty(); });
        b = queue.back();
        queue.pop_back();
      }
      work(*b);
      {
        std::lock_guard<std::mutex> lk(mu);
        --b->pending;
      }
      done.notify_all();
    }
  }
  std::mutex mu;
  std::condition_variable cv, done;
  int nworkers = 0;
  std::vector<Batch*> queue;
};

}  // namespace

extern "C" int pcc_stage_h2d(const void* const* srcs, const int64_t* nbytes, int32_t n_src, void* pinned, void* dev,
                             int64_t piece_bytes, int32_t threads, int32_t convert, pcc_stream_t s) {
  PCC_CHECK_ARG(n_src >= 0 && (n_src == 0 || (srcs && nbytes)) && pinned && dev);
  PCC_CHECK_ARG(piece_bytes >= 4096 && piece_bytes % 64 == 0 && threads >= 1 && threads <= 16);
  PCC_CHECK_ARG(convert == PCC_STAGE_COPY || convert == PCC_STAGE_F64_TO_BF16);
  Batch b;
  b.convert = convert;
  const int64_t shrink = convert ? 4 : 1;
  PCC_CUDA(cudaGetDevice(&b.device));
  b.stream = (cudaStream_t)s;
  int64_t pos = 0;
  for (int32_t i = 0; i < n_src; ++i) {
    PCC_CHECK_ARG(nbytes[i] >= 0 && (nbytes[i] == 0 || srcs[i]) && (!convert || nbytes[i] % 8 == 0));
    for (int64_t o = 0; o < nbytes[i]; o += piece_bytes) {
      const int64_t len = nbytes[i] - o < piece_bytes ? nbytes[i] - o : piece_bytes;
      b.pieces.push_back({(const uint8_t*)srcs[i] + o, (uint8_t*)pinned + pos + o / shrink, (uint8_t*)dev + pos + o / shrink, len,
                          len / shrink});
    }
    pos += nbytes[i] / shrink;
  }
  if (b.pieces.empty()) return PCC_OK;
  const int use = (int)b.pieces.size() < threads ? (int)b.pieces.size() : threads;
  if (Pool::get().run(b, use)) {
    pcc::set_error("pcc_stage_h2d: cudaMemcpyAsync failed (%s)", cudaGetErrorString(cudaGetLastError()));
    return PCC_E_CUDA;
  }
  return PCC_OK;
}
