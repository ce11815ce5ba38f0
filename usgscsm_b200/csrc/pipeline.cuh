// pipeline.cuh -- the chunked copy pipeline behind every HOST-pointer entry point.
//
// The arrays are cut into chunks, and each chunk goes through one of a few pipeline slots --
// H2D copies, the kernel(s), D2H copies, all on the slot's own stream -- so that while chunk c
// computes, chunk c+1 is copied in and chunk c-1 is copied out.  Copies are truly asynchronous
// when the caller's arrays are page-locked (cudaHostAlloc / cudaHostRegister / torch
// pin_memory).  Pageable arrays (what a numpy caller passes) are staged through the model's own
// page-locked ring by one copier thread per slot (`staged` below); the CUDA runtime's internal
// staging of pageable memory runs at a fraction of the link rate (profiles/r1_pageable_probe.txt).
#ifndef USGSCSM_B200_PIPELINE_CUH_
#define USGSCSM_B200_PIPELINE_CUH_

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "capi_common.h"

namespace ub200 {

struct HostIn {
  const void *p;      // host array
  int bytesPerPoint;  // 8 per double per point (24 for an array of csm::EcefCoord), 2 / 4 for DEM samples
  bool broadcast;     // one value for all points (height_stride = 0): a single element is copied
  // Rows of a larger host raster (a DEM window): rowPoints points per row, consecutive rows
  // rowPitchBytes apart.  0 = one contiguous array.  Chunks are then whole rows (granule).
  long rowPoints;
  long rowPitchBytes;
  HostIn(const double *ptr, int doublesPerPoint, bool bcast)
      : p(ptr), bytesPerPoint(doublesPerPoint * 8), broadcast(bcast), rowPoints(0), rowPitchBytes(0) {}
  HostIn(const void *ptr, int bytes, long rowPts, long pitchBytes)
      : p(ptr), bytesPerPoint(bytes), broadcast(false), rowPoints(rowPts), rowPitchBytes(pitchBytes) {}
};
struct HostOut {
  double *p;  // host array; nullptr = the caller does not want it (device pointer is null too)
  int perPoint;  // doubles per point
};

// Device views of one chunk inside a slot's scratch, handed to the launch callback.
struct Chunk {
  long offset;  // index of the chunk's first point in the caller's arrays
  long count;
  std::vector<double *> in, out, work;  // 16-byte aligned; `in` of a non-double input: cast back
  uint32_t *status;  // null when the caller passed no status array
  cudaStream_t stream;
};

// Set on the per-device worker threads of the _multi entry points: they already run one host
// thread per device, each keeps its chunk pipeline on that thread.
extern thread_local bool tlsSingleThreadPipeline;

bool pipelineThreads();  // USGSCSM_B200_PIPELINE_THREADS=0: pageable arrays are left to the CUDA runtime's staging
int stagedSlots();       // USGSCSM_B200_STAGE_THREADS (default 12): copier threads = staging slots
long stagedChunk();      // USGSCSM_B200_STAGE_CHUNK (default 256 Ki points)
bool hostPointerPageable(const void *p);

inline long roundUp16(long bytes) { return (bytes + 15) & ~15L; }
constexpr long kSmallCall = 4096;  // points at or below which a host-pointer call takes the single-copy path
constexpr long kZeroCopyCall = 64;  // ... and at or below which the kernels work on the page-locked block directly

// launch(chunk) enqueues the kernels of one chunk on chunk.stream and returns a cudaError_t.
// workPerPoint: extra device doubles per point the launch needs (AoS split / join).
// granule: chunks hold a whole number of granules (image rows for the dense-grid calls).
template <class Launch>
int runPipelined(usgscsm_b200_model *m, int64_t n, const std::vector<HostIn> &ins,
                 const std::vector<HostOut> &outs, uint32_t *status, int workPerPoint,
                 Launch &&launch, long granule = 1) {
  return guarded([&]() -> int {
    // the slots (scratch, streams, events) belong to the model: one host-pointer call at a time
    // per handle; a second thread calling on the same handle waits here (the reference's const
    // per-point calls are re-entrant, include/usgscsm/UsgsAstroLsSensorModel.h:910-914)
    std::lock_guard<std::mutex> oneCallPerModel(m->hostCallMutex);
    UB_CUDA(cudaSetDevice(m->device));
    // A handful of points (the scalar calls of a CSM application arrive here as batches of one):
    // latency is all that counts.  Everything goes through the model's page-locked block laid out
    // like the device scratch -- ONE copy up, the kernels, ONE copy down, on the calling thread.
    if (n <= kSmallCall && granule <= n) {
      const long cnt = (long)n;
      long inBytes = 0;
      for (const HostIn &a : ins) inBytes += roundUp16(cnt * a.bytesPerPoint);
      long total = inBytes;
      for (const HostOut &a : outs) total += roundUp16(cnt * a.perPoint * 8);
      total += (long)workPerPoint * roundUp16(cnt * 8) + roundUp16(cnt * 4);
      m->ensureScratch(0, (total + 7) / 8);
      m->ensureStage(0, total);
      PipelineSlot &s = m->slots[0];
      char *stage = static_cast<char *>(s.stage);
      // A few points: the kernels read and write the page-locked block itself (it is mapped into
      // the device's address space under unified addressing) -- no copy at all, a launch and a
      // synchronisation.  Beyond that the block is copied (latency-bound reads over the link cost
      // more than one DMA transfer).
      const bool zeroCopy = n <= kZeroCopyCall;
      char *base = zeroCopy ? stage : reinterpret_cast<char *>(s.scratch);
      Chunk k;
      k.offset = 0;
      k.count = cnt;
      k.stream = s.stream;
      long off = 0;
      for (const HostIn &a : ins) {
        k.in.push_back(reinterpret_cast<double *>(base + off));
        if (a.broadcast) {
          std::memcpy(stage + off, a.p, (size_t)a.bytesPerPoint);
        } else if (a.rowPoints > 0) {
          const long rows = cnt / a.rowPoints;
          const size_t width = (size_t)a.rowPoints * a.bytesPerPoint;
          for (long r = 0; r < rows; r++)
            std::memcpy(stage + off + r * width, static_cast<const char *>(a.p) + r * a.rowPitchBytes, width);
        } else {
          std::memcpy(stage + off, a.p, (size_t)cnt * a.bytesPerPoint);
        }
        off += roundUp16(cnt * a.bytesPerPoint);
      }
      for (const HostOut &a : outs) {
        k.out.push_back(a.p ? reinterpret_cast<double *>(base + off) : nullptr);
        off += roundUp16(cnt * a.perPoint * 8);
      }
      for (int w = 0; w < workPerPoint; w++, off += roundUp16(cnt * 8))
        k.work.push_back(reinterpret_cast<double *>(base + off));
      k.status = status ? reinterpret_cast<uint32_t *>(base + off) : nullptr;
      if (!zeroCopy && inBytes > 0)
        UB_CUDA(cudaMemcpyAsync(base, stage, (size_t)inBytes, cudaMemcpyHostToDevice, s.stream));
      UB_CUDA(launch(k));
      if (!zeroCopy)
        UB_CUDA(cudaMemcpyAsync(stage + inBytes, base + inBytes, (size_t)(total - inBytes), cudaMemcpyDeviceToHost,
                                s.stream));
      UB_CUDA(cudaStreamSynchronize(s.stream));
      for (size_t j = 0; j < outs.size(); j++)
        if (outs[j].p)
          std::memcpy(outs[j].p, stage + (reinterpret_cast<char *>(k.out[j]) - base),
                      (size_t)cnt * outs[j].perPoint * 8);
      if (status) std::memcpy(status, stage + (reinterpret_cast<char *>(k.status) - base), cnt * sizeof(uint32_t));
      return USGSCSM_B200_OK;
    }
    bool pageableIn = false, pageableOut = false;
    for (const HostIn &a : ins)
      if (!a.broadcast) {
        pageableIn = hostPointerPageable(a.p);
        break;
      }
    for (const HostOut &a : outs)
      if (a.p) {
        pageableOut = hostPointerPageable(a.p);
        break;
      }
    // Pageable arrays go through the model's own page-locked staging ring: more, smaller chunks in
    // flight, one host thread per slot, so that the host-side memcpy of one chunk overlaps the
    // link transfers and kernels of the others.
    const bool staged = (pageableIn || pageableOut) && pipelineThreads();
    const int nslots = staged ? stagedSlots() : pipelineSlots();
    const long wanted = staged ? stagedChunk() : pipelineChunk();
    long chunk = n < wanted ? (long)n : wanted;
    if (granule > 1) chunk = chunk < granule ? granule : chunk / granule * granule;
    long chunkBytes = roundUp16(chunk * 4);  // status words
    for (const HostIn &a : ins) chunkBytes += roundUp16(chunk * a.bytesPerPoint);
    for (const HostOut &a : outs) chunkBytes += roundUp16(chunk * a.perPoint * 8);
    chunkBytes += (long)workPerPoint * roundUp16(chunk * 8);
    const long nChunks = (n + chunk - 1) / chunk;

    // One chunk through slot si: copies in, kernels, copies out, all on the slot's stream.
    // stage != nullptr: the chunk's host side is the slot's page-locked staging buffer, laid out
    // like the device scratch (same offsets); the caller's arrays are copied to / from it by the
    // slot's host thread.
    auto runChunk = [&](long c, int si, char *stage) {
      const long off = c * chunk;
      const long cnt = (n - off) < chunk ? (long)(n - off) : chunk;
      PipelineSlot &s = m->slots[si];
      Chunk k;
      k.offset = off;
      k.count = cnt;
      k.stream = s.stream;
      char *base = reinterpret_cast<char *>(s.scratch);
      char *next = base;
      for (const HostIn &a : ins) {
        k.in.push_back(reinterpret_cast<double *>(next));
        if (a.broadcast) {
          UB_CUDA(cudaMemcpyAsync(next, a.p, a.bytesPerPoint, cudaMemcpyHostToDevice, s.stream));
        } else if (a.rowPoints > 0) {
          const long rows = cnt / a.rowPoints, row0 = off / a.rowPoints;
          const size_t width = (size_t)a.rowPoints * a.bytesPerPoint;
          const char *src = static_cast<const char *>(a.p) + row0 * a.rowPitchBytes;
          if (stage && pageableIn) {
            char *h = stage + (next - base);
            for (long r = 0; r < rows; r++) std::memcpy(h + r * width, src + r * a.rowPitchBytes, width);
            UB_CUDA(cudaMemcpyAsync(next, h, rows * width, cudaMemcpyHostToDevice, s.stream));
          } else {
            UB_CUDA(cudaMemcpy2DAsync(next, width, src, (size_t)a.rowPitchBytes, width, (size_t)rows,
                                      cudaMemcpyHostToDevice, s.stream));
          }
        } else {
          const char *src = static_cast<const char *>(a.p) + off * a.bytesPerPoint;
          const size_t bytes = (size_t)cnt * a.bytesPerPoint;
          if (stage && pageableIn) {
            char *h = stage + (next - base);
            std::memcpy(h, src, bytes);
            UB_CUDA(cudaMemcpyAsync(next, h, bytes, cudaMemcpyHostToDevice, s.stream));
          } else {
            UB_CUDA(cudaMemcpyAsync(next, src, bytes, cudaMemcpyHostToDevice, s.stream));
          }
        }
        next += roundUp16(chunk * a.bytesPerPoint);
      }
      for (const HostOut &a : outs) {
        k.out.push_back(a.p ? reinterpret_cast<double *>(next) : nullptr);
        next += roundUp16(chunk * a.perPoint * 8);
      }
      for (int w = 0; w < workPerPoint; w++, next += roundUp16(chunk * 8))
        k.work.push_back(reinterpret_cast<double *>(next));
      k.status = status ? reinterpret_cast<uint32_t *>(next) : nullptr;
      UB_CUDA(launch(k));
      const bool viaStage = stage && pageableOut;
      for (size_t j = 0; j < outs.size(); j++)
        if (outs[j].p) {
          const size_t bytes = (size_t)cnt * outs[j].perPoint * 8;
          char *dev = reinterpret_cast<char *>(k.out[j]);
          void *dst = viaStage ? static_cast<void *>(stage + (dev - base))
                               : static_cast<void *>(outs[j].p + off * outs[j].perPoint);
          UB_CUDA(cudaMemcpyAsync(dst, dev, bytes, cudaMemcpyDeviceToHost, s.stream));
        }
      if (status) {
        char *dev = reinterpret_cast<char *>(k.status);
        void *dst = viaStage ? static_cast<void *>(stage + (dev - base)) : static_cast<void *>(status + off);
        UB_CUDA(cudaMemcpyAsync(dst, dev, cnt * sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
      }
      if (viaStage) {
        // results land in the page-locked buffer; hand them to the caller's arrays
        UB_CUDA(cudaStreamSynchronize(s.stream));
        for (size_t j = 0; j < outs.size(); j++)
          if (outs[j].p)
            std::memcpy(outs[j].p + off * outs[j].perPoint,
                        stage + (reinterpret_cast<char *>(k.out[j]) - base),
                        (size_t)cnt * outs[j].perPoint * 8);
        if (status)
          std::memcpy(status + off, stage + (reinterpret_cast<char *>(k.status) - base),
                      cnt * sizeof(uint32_t));
      }
    };
    const int used = (int)(nChunks < nslots ? nChunks : nslots);
    for (int si = 0; si < used; si++) m->ensureScratch(si, (chunkBytes + 7) / 8);

    if (staged) {
      for (int si = 0; si < used; si++) m->ensureStage(si, chunkBytes);
      std::vector<int> codes((size_t)used, USGSCSM_B200_OK);
      std::vector<std::string> messages((size_t)used);
      std::vector<std::thread> workers;
      for (int t = 0; t < used; t++)
        workers.emplace_back([&, t] {
          codes[(size_t)t] = guarded([&]() -> int {
            UB_CUDA(cudaSetDevice(m->device));
            for (long c = t; c < nChunks; c += used) {
              runChunk(c, t, static_cast<char *>(m->slots[t].stage));
              UB_CUDA(cudaStreamSynchronize(m->slots[t].stream));
            }
            return USGSCSM_B200_OK;
          });
          if (codes[(size_t)t] != USGSCSM_B200_OK) messages[(size_t)t] = lastError();
        });
      for (std::thread &w : workers) w.join();
      for (int t = 0; t < used; t++)
        if (codes[(size_t)t] != USGSCSM_B200_OK) {
          lastError() = messages[(size_t)t];
          return codes[(size_t)t];
        }
      return USGSCSM_B200_OK;
    }

    for (long c = 0; c < nChunks; c++) {
      const int si = (int)(c % used);
      PipelineSlot &s = m->slots[si];
      if (c >= used) UB_CUDA(cudaEventSynchronize(s.done));  // the slot's previous chunk
      runChunk(c, si, nullptr);
      UB_CUDA(cudaEventRecord(s.done, s.stream));
    }
    for (int si = 0; si < used; si++) UB_CUDA(cudaStreamSynchronize(m->slots[si].stream));
    return USGSCSM_B200_OK;
  });
}

inline bool badCall(usgscsm_b200_model *m, int64_t n, int &rc) {
  if (!m || n < 0) {
    rc = fail(USGSCSM_B200_EINVAL, "bad argument");
    return true;
  }
  if (n == 0) {
    rc = USGSCSM_B200_OK;
    return true;
  }
  return false;
}

}  // namespace ub200
#endif
