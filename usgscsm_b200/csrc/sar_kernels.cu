// sar_kernels.cu -- SAR groundToImage / imageToGround, one thread per point, FP64.
//
// Replaces UsgsAstroSarSensorModel::groundToImage (src/UsgsAstroSarSensorModel.cpp:541-593:
// zero-Doppler time by brentRoot src/Utilities.cpp:545-622 over dopplerShift SAR:595-632,
// slantRange SAR:634-656, slantRangeToGroundRange SAR:658-686 with polynomialRoot
// U:771-791 over the time-interpolated range polynomial getRangeCoefficients SAR:1936-1972)
// and ::imageToGround (SAR:775-861).  Exceptions of the scalar code (no Doppler bracket,
// Newton failure) become USGSCSM_B200_ST_ERROR_UNKNOWN with NaN outputs (SAR:569-575).
//
// The ephemeris tables (positions, velocities: ~1200 nodes each for an orbital SAR image)
// and the range-coefficient table are staged once per persistent block into shared memory.
#include <cstdlib>

#include "kernels.h"
#include "device_math.cuh"
#include "work_queue.cuh"

#ifndef UB_SAR_G2I_MINB
#define UB_SAR_G2I_MINB 3
#endif

namespace ub200 {

namespace {

constexpr int kThreads = 256;
constexpr size_t kSarMaxStagedBytes = 100 * 1024;  // 2 resident blocks per SM

// Where a kernel reads the tables from: the device arrays of the packed state (rows of 3 doubles),
// or the block's copy in shared memory, rows padded to 4 doubles on 16-byte boundaries so that a
// node is two 128-bit loads (unordered points make every lane read its own nodes: fewer, wider
// shared-memory transactions).
// LAYOUT 0: rows of 3 doubles (the device arrays, or a plain copy in shared memory); 1: the
// padded copy; 2: positions and velocities INTERLEAVED in shared memory, one row of 6 doubles
// (px, py, pz, vx, vy, vz) per node on a 16-byte boundary -- the Doppler function needs both
// vectors at the same 8 nodes, and three 128-bit loads per node instead of six 64-bit ones halve
// the shared-memory instructions of unordered points (every lane reads its own nodes, so the
// search is bound by shared-memory wavefronts, not arithmetic).  `positions` then points at the
// interleaved table and `velocities` is unused.
constexpr int kSarPlain = 0, kSarPadded = 1, kSarInterleaved = 2;
template <int LAYOUT>
struct SarTablesT {
  static constexpr bool kPadded = LAYOUT == kSarPadded;
  static constexpr bool kInterleaved = LAYOUT == kSarInterleaved;
  const double *positions;
  const double *velocities;
  const double *scale;
};
using SarTables = SarTablesT<kSarPlain>;

template <class TAB>
UB_DEV void sarApply3(const TAB &, const double *table, const LagrangeWeights &w, double out[3]) {
  if (TAB::kPadded)
    lagrangeApplyPadded<3>(table, w, out);
  else
    lagrangeApply<3>(table, w, out);
}
// (the interleaved layout takes the same room as the plain one: 6 doubles per node)
__host__ __device__ inline size_t sarTableBytes(const SarState &S, int layout = kSarPlain) {
  const size_t row = layout == kSarPadded ? 4 : 3;
  return ((size_t)S.positions.numNodes * row + (size_t)S.velocities.numNodes * row +
          (size_t)S.scale.numNodes * 4) *
         sizeof(double);
}

// the interleaved layout needs the two tables on the same nodes (they share t0 / dt in the state)
__host__ __device__ inline bool sarInterleavable(const SarState &S) {
  return S.positions.numNodes > 1 && S.positions.numNodes == S.velocities.numNodes;
}

// STAGE: copy the tables into `smem` in layout LAYOUT (the launcher has checked
// sarInterleavable for layout 2).
template <bool STAGE, int LAYOUT = kSarPlain>
__device__ inline SarTablesT<STAGE ? LAYOUT : kSarPlain> stageSarTables(const SarState &S, double *smem) {
  SarTablesT<STAGE ? LAYOUT : kSarPlain> t{S.positions.values, S.velocities.values, S.scale.values};
  if (!STAGE) return t;
  constexpr bool PAD = LAYOUT == kSarPadded;
  constexpr int ROW = PAD ? 4 : 3;
  const int np = S.positions.numNodes * ROW, nv = S.velocities.numNodes * ROW,
            ns = S.scale.numNodes * 4;
  double *sp = smem, *sv = sp + np, *ss = sv + nv;
  if (LAYOUT == kSarInterleaved) {
    for (int i = threadIdx.x; i < np + nv; i += blockDim.x) {
      const int node = i / 6, j = i - node * 6;
      sp[i] = j < 3 ? __ldg(S.positions.values + 3 * node + j) : __ldg(S.velocities.values + 3 * node + j - 3);
    }
  } else if (PAD) {
    for (int i = threadIdx.x; i < np; i += blockDim.x)
      sp[i] = (i & 3) < 3 ? __ldg(S.positions.values + 3 * (i >> 2) + (i & 3)) : 0.0;
    for (int i = threadIdx.x; i < nv; i += blockDim.x)
      sv[i] = (i & 3) < 3 ? __ldg(S.velocities.values + 3 * (i >> 2) + (i & 3)) : 0.0;
  } else {
    for (int i = threadIdx.x; i < np; i += blockDim.x) sp[i] = __ldg(S.positions.values + i);
    for (int i = threadIdx.x; i < nv; i += blockDim.x) sv[i] = __ldg(S.velocities.values + i);
  }
  for (int i = threadIdx.x; i < ns; i += blockDim.x) ss[i] = __ldg(S.scale.values + i);
  __syncthreads();
  t.positions = sp;
  t.velocities = sv;
  t.scale = ss;
  return t;
}

// getAdjustedSensorPosition SAR:1196-1218 and getAdjustedSensorVelocity SAR:1266-1288 with
// adj = m_noAdjustments.  Both tables share t0 / dt; when they also have the same node
// count the Lagrange weights are identical and evaluated once.
template <class TAB>
UB_DEV void sarPosVel(const SarState &S, const TAB &T, double time, double pos[3],
                      double vel[3]) {
  LagrangeWeights w;
  if (TAB::kInterleaved) {
    lagrangeWeights(S.positions, S.lag, time, w);
    double pv[6];
    lagrangeApplyInterleaved6(T.positions, w, pv);
    pos[0] = pv[0] + S.params[0];
    pos[1] = pv[1] + S.params[1];
    pos[2] = pv[2] + S.params[2];
    vel[0] = pv[3] + S.params[3];
    vel[1] = pv[4] + S.params[4];
    vel[2] = pv[5] + S.params[5];
    return;
  }
  bool haveW = false;
  if (S.positions.numNodes > 1) {
    lagrangeWeights(S.positions, S.lag, time, w);
    haveW = true;
    double p[3];
    sarApply3(T, T.positions, w, p);
    pos[0] = p[0] + S.params[0];
    pos[1] = p[1] + S.params[1];
    pos[2] = p[2] + S.params[2];
  } else {
    pos[0] = T.positions[0] + S.params[0];
    pos[1] = T.positions[1] + S.params[1];
    pos[2] = T.positions[2] + S.params[2];
  }
  if (S.velocities.numNodes > 1) {
    if (!haveW || S.velocities.numNodes != S.positions.numNodes)
      lagrangeWeights(S.velocities, S.lag, time, w);
    double v[3];
    sarApply3(T, T.velocities, w, v);
    vel[0] = v[0] + S.params[3];
    vel[1] = v[1] + S.params[4];
    vel[2] = v[2] + S.params[5];
  } else {
    vel[0] = T.velocities[0] + S.params[3];
    vel[1] = T.velocities[1] + S.params[4];
    vel[2] = T.velocities[2] + S.params[5];
  }
}

// The Doppler-shift function of dopplerShift SAR:603-617; also returns the slant range.
template <class TAB>
UB_DEV double sarDoppler(const SarState &S, const TAB &T, double time, double gx, double gy,
                         double gz, double &slantRange) {
  double pos[3], vel[3];
  sarPosVel(S, T, time, pos, vel);
  double lx = pos[0] - gx, ly = pos[1] - gy, lz = pos[2] - gz;
  slantRange = sqrt(lx * lx + ly * ly + lz * lz);
  return -2.0 * (lx * vel[0] + ly * vel[1] + lz * vel[2]) / (slantRange * S.wavelength);
}

// getRangeCoefficients SAR:1936-1972
template <class TAB>
UB_DEV void sarRangeCoefficients(const SarState &S, const TAB &T, double time, double c[4]) {
  if (S.scale.numNodes > 1) {
    LagrangeWeights w;
    lagrangeWeights(S.scale, S.lag, time, w);
    lagrangeApply<4>(T.scale, w, c);
  } else {
    c[0] = T.scale[0];
    c[1] = T.scale[1];
    c[2] = T.scale[2];
    c[3] = T.scale[3];
  }
}

// ---- SAR::groundToImage SAR:541-593 as a resumable search ----------------------------------------
// The zero-Doppler search (brentRoot U:545-622 over dopplerShift SAR:595-632) takes 10 to 32
// function evaluations depending on the point, so points that share a warp finish at different
// times.  The search is therefore written as a state that advances by ONE evaluation per call:
// the kernel below keeps every lane busy by handing a finished lane the next point at once.
struct SarSearch {
  double gx, gy, gz;
  BrentSearch r;
  double slant;  // slant range of the last evaluation
};

UB_DEV void sarSearchBegin(const SarState &S, double gx, double gy, double gz, SarSearch &q) {
  const double timePadding = S.exposureDuration * S.nLines * 0.25;
  q.gx = gx;
  q.gy = gy;
  q.gz = gz;
  q.slant = 0.0;
  brentBegin(S.startingEphemerisTime - timePadding, S.endingEphemerisTime + timePadding, q.r);
}

// One evaluation of the Doppler function and the bookkeeping that follows it in brentRoot.
// True when the search has ended (q.r.t, q.slant, q.r.failed, q.r.status are final).
template <class TAB>
UB_DEV bool sarSearchStep(const SarState &S, const TAB &T, double timeTolerance, SarSearch &q) {
  return brentStep([&](double t) { return sarDoppler(S, T, t, q.gx, q.gy, q.gz, q.slant); },
                   timeTolerance, q.r);
}

// After the search: time = the last abscissa evaluated; slantRange(groundPt, time) SAR:634-656 is
// the norm that evaluation computed; ground range by polynomialRoot U:771-791 (maxIterations 30).
template <class TAB>
UB_DEV void sarSearchFinish(const SarState &S, const TAB &T, double time, double slant, bool failed,
                            double desiredPrecision, double &outLine, double &outSamp,
                            uint32_t &status) {
  if (!failed) {
    double c[4];
    sarRangeCoefficients(S, T, time, c);
    double guess = 0.5 * S.nSamples * S.scaledPixelWidth;
    const double groundTolerance = S.scaledPixelWidth * desiredPrecision / 2.0;
    c[0] -= slant;
    double root;
    const bool converged = polynomialRoot3(c, guess, groundTolerance, root);
    if (converged) {
      outLine = divideBy(time - S.startingEphemerisTime, S.exposure) + 0.5;
      outSamp = divideBy(root, S.pixelWidth) + 0.5;
      return;
    }
  }
  status |= USGSCSM_B200_ST_ERROR_UNKNOWN;
  outLine = nan("");
  outSamp = nan("");
}

// One point, start to end (the callers that run several searches per point).
template <class TAB>
UB_DEV void sarGroundToImagePoint(const SarState &S, const TAB &T, double gx, double gy,
                                  double gz, double desiredPrecision, double &outLine,
                                  double &outSamp, uint32_t &status) {
  const double timeTolerance = S.exposureDuration * desiredPrecision / 2.0;
  SarSearch q;
  sarSearchBegin(S, gx, gy, gz, q);
  while (!sarSearchStep(S, T, timeTolerance, q)) {
  }
  status |= q.r.status | (q.r.evals << USGSCSM_B200_ST_EVALS_SHIFT);
  sarSearchFinish(S, T, q.r.t, q.slant, q.r.failed, desiredPrecision, outLine, outSamp, status);
}

// One point of SAR::imageToGround SAR:775-861.
template <class TAB>
UB_DEV void sarImageToGroundPoint(const SarState &S, const TAB &T, double line, double samp,
                                  double height, double desiredPrecision, double &x, double &y,
                                  double &z, double &achieved, uint32_t &status) {
  double time = S.startingEphemerisTime + (line - 0.5) * S.exposureDuration;
  double groundRange = (samp - 0.5) * S.scaledPixelWidth;
  double c[4];
  sarRangeCoefficients(S, T, time, c);
  double slantRange = poly3(c, groundRange);

  double pos[3], vel[3];
  sarPosVel(S, T, time, pos, vel);
  double positionMag = sqrt(pos[0] * pos[0] + pos[1] * pos[1] + pos[2] * pos[2]);
  // vHat = normalized(velocity) = 1.0 / norm * vec  (U:899-901, 972-974)
  double s = 1.0 / sqrt(vel[0] * vel[0] + vel[1] * vel[1] + vel[2] * vel[2]);
  double vHat[3] = {s * vel[0], s * vel[1], s * vel[2]};
  // tHat = normalized(rejection(position, vHat))
  double pr = (pos[0] * vHat[0] + pos[1] * vHat[1] + pos[2] * vHat[2]) /
              (vHat[0] * vHat[0] + vHat[1] * vHat[1] + vHat[2] * vHat[2]);
  double rej[3] = {pos[0] - pr * vHat[0], pos[1] - pr * vHat[1], pos[2] - pr * vHat[2]};
  s = 1.0 / sqrt(rej[0] * rej[0] + rej[1] * rej[1] + rej[2] * rej[2]);
  double tHat[3] = {s * rej[0], s * rej[1], s * rej[2]};
  double uHat[3] = {vHat[1] * tHat[2] - vHat[2] * tHat[1], vHat[2] * tHat[0] - vHat[0] * tHat[2],
                    vHat[0] * tHat[1] - vHat[1] * tHat[0]};
  double nadirComp = pos[0] * tHat[0] + pos[1] * tHat[1] + pos[2] * tHat[2];

  double pointRadius = S.ell.majorAxis + height;
  double pointHeight = 0.0;
  double gvx = 0.0, gvy = 0.0, gvz = 0.0;
  for (int it = 0; it < 10; it++) {
    double radiusSqr = pointRadius * pointRadius;
    double alpha =
        (radiusSqr - slantRange * slantRange - positionMag * positionMag) / (2 * nadirComp);
    double beta = sqrt(slantRange * slantRange - alpha * alpha);
    if (S.lookDirection == 0) beta *= -1;
    gvx = alpha * tHat[0] + beta * uHat[0] + pos[0];
    gvy = alpha * tHat[1] + beta * uHat[1] + pos[1];
    gvz = alpha * tHat[2] + beta * uHat[2] + pos[2];
    pointHeight = ellipsoidElevation(gvx, gvy, gvz, S.ell, 0.001);
    pointRadius -= (pointHeight - height);
    if (fabs(pointHeight - height) <= desiredPrecision) break;
  }
  achieved = fabs(pointHeight - height);
  x = gvx;
  y = gvy;
  z = gvz;
  if (!finite3(x, y, z)) status |= USGSCSM_B200_ST_NONFINITE;
}

// groundToImage for a batch.  Every lane runs its own point's search one evaluation per loop
// pass; a lane whose search has ended parks the result (time, slant range) in the warp's queue
// in shared memory and takes the next point from the work queue straight away, so the warp
// never idles behind its slowest point (mean 26 evaluations per point, 32 for the slowest: a
// warp that waits runs every search to 32).  When the queue holds a warp's worth of results
// all 32 lanes finish them together (range polynomial, stores).
constexpr int kSarQueue = 64;  // entries per warp: < 32 left after a flush + at most 32 pushed
struct SarWarpQueue {
  double time[kSarQueue];
  double slant[kSarQueue];
  long index[kSarQueue];
  uint32_t status[kSarQueue];  // status bits and the evaluation count; bit 31: no bracket
};

template <bool STAGE, int LAYOUT, int MINB>
__global__ void __launch_bounds__(kThreads, MINB)
    sarGroundToImageRefillKernel(const __grid_constant__ SarState S, long n, const double *__restrict__ x,
                           const double *__restrict__ y, const double *__restrict__ z, double prec,
                           double *__restrict__ line, double *__restrict__ samp,
                           double *__restrict__ achieved, uint32_t *__restrict__ status,
                           unsigned long long *__restrict__ next, int grab) {
  extern __shared__ __align__(16) double smem[];
  SarWarpQueue *queues = reinterpret_cast<SarWarpQueue *>(smem);
  SarWarpQueue &wq = queues[threadIdx.x >> 5];
  const SarTablesT<STAGE ? LAYOUT : kSarPlain> T =
      stageSarTables<STAGE, LAYOUT>(S, smem + (kThreads / 32) * sizeof(SarWarpQueue) / sizeof(double));
  const unsigned lane = threadIdx.x & 31u;
  const unsigned below = (1u << lane) - 1u;
  const double timeTolerance = S.exposureDuration * prec / 2.0;

  long chunkNext = 0, chunkEnd = 0;  // this warp's current grab of the work queue (warp-uniform)
  bool drained = false;              // the work queue is empty
  int queued = 0;                    // results waiting in wq (warp-uniform)
  bool busy = false;                 // this lane has a search in flight
  long point = 0;
  SarSearch q;

  for (;;) {
    // idle lanes take the next points
    unsigned idle = __ballot_sync(0xffffffffu, !busy);
    while (idle && !drained) {
      if (chunkNext >= chunkEnd) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(next, (unsigned long long)grab);
        base = __shfl_sync(0xffffffffu, base, 0);
        if ((long)base >= n) {
          drained = true;
          break;
        }
        chunkNext = (long)base;
        chunkEnd = chunkNext + grab < n ? chunkNext + grab : n;
      }
      const long avail = chunkEnd - chunkNext;
      const int rank = __popc(idle & below);
      if (!busy && rank < avail) {
        point = chunkNext + rank;
        sarSearchBegin(S, __ldcs(x + point), __ldcs(y + point), __ldcs(z + point), q);
        busy = true;
      }
      const int wanted = __popc(idle);
      chunkNext += wanted < avail ? wanted : avail;
      idle = __ballot_sync(0xffffffffu, !busy);
    }
    const unsigned active = __ballot_sync(0xffffffffu, busy);
    if (active) {
      bool ended = false;
      if (busy) ended = sarSearchStep(S, T, timeTolerance, q);
      const unsigned done = __ballot_sync(0xffffffffu, ended);
      if (ended) {
        const int slot = queued + __popc(done & below);
        wq.time[slot] = q.r.t;
        wq.slant[slot] = q.slant;
        wq.index[slot] = point;
        wq.status[slot] =
            q.r.status | (q.r.evals << USGSCSM_B200_ST_EVALS_SHIFT) | (q.r.failed ? 0x80000000u : 0u);
        busy = false;
      }
      queued += __popc(done);
      __syncwarp();
    }
    // finish a warp's worth of results together; at the very end, what is left
    if (queued >= 32 || (!active && queued > 0)) {
      const int take = queued < 32 ? queued : 32;
      if ((int)lane < take) {
        const long i = wq.index[lane];
        uint32_t st = wq.status[lane];
        const bool failed = (st & 0x80000000u) != 0;
        st &= 0x7fffffffu;
        double l, s;
        sarSearchFinish(S, T, wq.time[lane], wq.slant[lane], failed, prec, l, s, st);
        if (!(isfinite(l) && isfinite(s))) st |= USGSCSM_B200_ST_NONFINITE;
        // getValidImageRange SAR:1066-1072: (0,0) .. (nLines, nSamples)
        if (!(l >= 0.0 && l <= 0.0 + S.nLines && s >= 0.0 && s <= 0.0 + S.nSamples))
          st |= USGSCSM_B200_ST_OUT_OF_IMAGE;
        __stcs(line + i, l);
        __stcs(samp + i, s);
        if (achieved) __stcs(achieved + i, nan(""));  // SAR groundToImage never writes achievedPrecision
        if (status) __stcs(status + i, st);
      }
      __syncwarp();
      const int left = queued - take;
      double mt = 0.0, ms = 0.0;
      long mi = 0;
      uint32_t mst = 0;
      if ((int)lane < left) {
        mt = wq.time[32 + lane];
        ms = wq.slant[32 + lane];
        mi = wq.index[32 + lane];
        mst = wq.status[32 + lane];
      }
      __syncwarp();
      if ((int)lane < left) {
        wq.time[lane] = mt;
        wq.slant[lane] = ms;
        wq.index[lane] = mi;
        wq.status[lane] = mst;
      }
      __syncwarp();
      queued = left;
    }
    if (!active && drained && queued == 0) break;
  }
}

// The plain form: a warp takes 32 points and runs their searches in lockstep.  Best when
// neighbouring points need the same number of evaluations (dense grids: 30.4 of 32 lanes active
// on average) -- and with the fewest registers.
template <bool STAGE, int LAYOUT, int MINB>
__global__ void __launch_bounds__(kThreads, MINB)
    sarGroundToImageKernel(const __grid_constant__ SarState S, long n, const double *__restrict__ x,
                           const double *__restrict__ y, const double *__restrict__ z, double prec,
                           double *__restrict__ line, double *__restrict__ samp,
                           double *__restrict__ achieved, uint32_t *__restrict__ status,
                           unsigned long long *__restrict__ next, int grab) {
  extern __shared__ __align__(16) double smem[];
  const SarTablesT<STAGE ? LAYOUT : kSarPlain> T = stageSarTables<STAGE, LAYOUT>(S, smem);
  forEachPointDynamic(n, next, grab, [&](long i) {
    double gx = __ldcs(x + i), gy = __ldcs(y + i), gz = __ldcs(z + i);
    double l, s;
    uint32_t st = 0;
    sarGroundToImagePoint(S, T, gx, gy, gz, prec, l, s, st);
    if (!(isfinite(l) && isfinite(s))) st |= USGSCSM_B200_ST_NONFINITE;
    // getValidImageRange SAR:1066-1072: (0,0) .. (nLines, nSamples)
    if (!(l >= 0.0 && l <= 0.0 + S.nLines && s >= 0.0 && s <= 0.0 + S.nSamples))
      st |= USGSCSM_B200_ST_OUT_OF_IMAGE;
    __stcs(line + i, l);
    __stcs(samp + i, s);
    if (achieved) __stcs(achieved + i, nan(""));  // SAR groundToImage never writes achievedPrecision
    if (status) __stcs(status + i, st);
  });
}

template <bool STAGE, bool GRID>
__global__ void __launch_bounds__(kThreads)
    sarImageToGroundKernel(const __grid_constant__ SarState S, long n, const double *__restrict__ line,
                           const double *__restrict__ samp, GridSpec g,
                           const double *__restrict__ height, int heightStride, double prec,
                           double *__restrict__ x, double *__restrict__ y, double *__restrict__ z,
                           double *__restrict__ achieved, uint32_t *__restrict__ status,
                           unsigned long long *__restrict__ next, int grab) {
  extern __shared__ __align__(16) double smem[];
  const SarTables T = stageSarTables<STAGE>(S, smem);
  forEachPointDynamic(n, next, grab, [&](long i) {
    double l, s;
    if (GRID) {
      long r = i / g.nSamps;
      long c = i - r * g.nSamps;
      l = g.line0 + (double)(g.row0 + r) * g.dline;
      s = g.samp0 + (double)c * g.dsamp;
    } else {
      l = __ldcs(line + i);
      s = __ldcs(samp + i);
    }
    double h = heightStride ? __ldcs(height + i) : __ldg(height);
    double px, py, pz, a;
    uint32_t st = 0;
    sarImageToGroundPoint(S, T, l, s, h, prec, px, py, pz, a, st);
    __stcs(x + i, px);
    __stcs(y + i, py);
    __stcs(z + i, pz);
    if (achieved) __stcs(achieved + i, a);
    if (status) __stcs(status + i, st);
  });
}

// computeGroundPartials SAR:1384-1406: delta = m_scaledPixelWidth.
template <bool STAGE>
__global__ void __launch_bounds__(kThreads)
    sarGroundPartialsKernel(const __grid_constant__ SarState S, long n, const double *__restrict__ x,
                            const double *__restrict__ y, const double *__restrict__ z,
                            double *__restrict__ partials, uint32_t *__restrict__ status,
                            unsigned long long *__restrict__ next, int grab) {
  extern __shared__ __align__(16) double smem[];
  const SarTables T = stageSarTables<STAGE>(S, smem);
  forEachPointDynamic(n, next, grab, [&](long i) {
    const double g[3] = {__ldcs(x + i), __ldcs(y + i), __ldcs(z + i)};
    const double delta = S.scaledPixelWidth;
    double l[4], s[4];
    uint32_t st = 0;
#pragma unroll 1
    for (int k = 0; k < 4; k++) {
      double q[3] = {g[0], g[1], g[2]};
      if (k) q[k - 1] = g[k - 1] + delta;
      uint32_t sk = 0;
      sarGroundToImagePoint(S, T, q[0], q[1], q[2], 0.001, l[k], s[k], sk);
      st |= sk;
    }
    double p[6];
    p[0] = (l[1] - l[0]) / delta;
    p[3] = (s[1] - s[0]) / delta;
    p[1] = (l[2] - l[0]) / delta;
    p[4] = (s[2] - s[0]) / delta;
    p[2] = (l[3] - l[0]) / delta;
    p[5] = (s[3] - s[0]) / delta;
    storePartials(partials, status, i, p, st);
  });
}

// imageToRemoteImagingLocus SAR:1009-1030: point = imageToGround at height 0, direction =
// normalized(ground - getSensorPosition(imagePt)) with normalized() = (1 / norm) * v (U:972-974);
// achievedPrecision and warnings are those of the imageToGround call.
template <bool STAGE>
__global__ void __launch_bounds__(kThreads)
    sarRemoteLocusKernel(const __grid_constant__ SarState S, long n, const double *__restrict__ line,
                         const double *__restrict__ samp, double prec, LocusOut o) {
  extern __shared__ __align__(16) double smem[];
  const SarTables T = stageSarTables<STAGE>(S, smem);
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double l = __ldcs(line + i), s = __ldcs(samp + i);
    double time = S.startingEphemerisTime + (l - 0.5) * S.exposureDuration;
    double sensor[3], vel[3];
    sarPosVel(S, T, time, sensor, vel);
    double g[3], a;
    uint32_t st = 0;
    sarImageToGroundPoint(S, T, l, s, 0.0, prec, g[0], g[1], g[2], a, st);
    double v[3] = {g[0] - sensor[0], g[1] - sensor[1], g[2] - sensor[2]};
    double inv = 1.0 / sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    double dir[3] = {inv * v[0], inv * v[1], inv * v[2]};
    if (!finite3(dir[0], dir[1], dir[2])) st |= USGSCSM_B200_ST_NONFINITE;
    storeLocus(o, i, g, dir, a, st);
  }
}

// imageToProximateImagingLocus SAR:959-993: the point at the pixel's slant range from the
// spacecraft, in the direction of the ground point's rejection from the velocity; direction =
// the tangent of the range circle there.  achievedPrecision = 0, no warnings.
template <bool STAGE>
__global__ void __launch_bounds__(kThreads)
    sarProximateLocusKernel(const __grid_constant__ SarState S, long n, const double *__restrict__ line,
                            const double *__restrict__ samp, const double *__restrict__ x,
                            const double *__restrict__ y, const double *__restrict__ z, LocusOut o) {
  extern __shared__ __align__(16) double smem[];
  const SarTables T = stageSarTables<STAGE>(S, smem);
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double time = S.startingEphemerisTime + (__ldcs(line + i) - 0.5) * S.exposureDuration;
    double groundRange = (__ldcs(samp + i) - 0.5) * S.scaledPixelWidth;
    double c[4];
    sarRangeCoefficients(S, T, time, c);
    double slantRange = poly3(c, groundRange);
    double p[3], v[3];
    sarPosVel(S, T, time, p, v);
    // lookVec = normalized(rejection(ground - position, velocity)), U:972-1001
    double w[3] = {__ldcs(x + i) - p[0], __ldcs(y + i) - p[1], __ldcs(z + i) - p[2]};
    double pr = (w[0] * v[0] + w[1] * v[1] + w[2] * v[2]) / (v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    double rej[3] = {w[0] - pr * v[0], w[1] - pr * v[1], w[2] - pr * v[2]};
    double inv = 1.0 / sqrt(rej[0] * rej[0] + rej[1] * rej[1] + rej[2] * rej[2]);
    double look[3] = {inv * rej[0], inv * rej[1], inv * rej[2]};
    double pt[3] = {p[0] + slantRange * look[0], p[1] + slantRange * look[1],
                    p[2] + slantRange * look[2]};
    double t[3];
    if (S.lookDirection == 0) {  // LEFT: cross(velocity, lookVec)
      t[0] = v[1] * look[2] - v[2] * look[1];
      t[1] = v[2] * look[0] - v[0] * look[2];
      t[2] = v[0] * look[1] - v[1] * look[0];
    } else {  // cross(lookVec, velocity)
      t[0] = look[1] * v[2] - look[2] * v[1];
      t[1] = look[2] * v[0] - look[0] * v[2];
      t[2] = look[0] * v[1] - look[1] * v[0];
    }
    inv = 1.0 / sqrt(t[0] * t[0] + t[1] * t[1] + t[2] * t[2]);
    double dir[3] = {inv * t[0], inv * t[1], inv * t[2]};
    uint32_t st = 0;
    if (!(finite3(pt[0], pt[1], pt[2]) && finite3(dir[0], dir[1], dir[2])))
      st |= USGSCSM_B200_ST_NONFINITE;
    storeLocus(o, i, pt, dir, 0.0, st);
  }
}

LaunchShape sarShape(const SarState &S, long n, int smCount, size_t &smemBytes, bool &stage,
                     size_t extraSmem = 0, int maxPerSm = kBlocksPerSm, int layout = kSarPlain) {
  size_t bytes = sarTableBytes(S, layout);
  stage = bytes <= kSarMaxStagedBytes;
  smemBytes = (stage ? bytes : 0) + extraSmem;
  int perSm = maxPerSm;
  if (smemBytes > 0) {
    size_t fit = (size_t)(220 * 1024) / (smemBytes + 1024);
    if (fit < (size_t)perSm) perSm = (int)(fit < 1 ? 1 : fit);
  }
  LaunchShape sh;
  sh.threads = kThreads;
  long blocksNeeded = (n + kThreads - 1) / kThreads;
  long persistent = (long)smCount * perSm;
  sh.blocks = (int)(blocksNeeded < persistent ? blocksNeeded : persistent);
  if (sh.blocks < 1) sh.blocks = 1;
  return sh;
}

int sarEnv(const char *name, int def) {
  const char *v = getenv(name);
  return v && *v ? atoi(v) : def;
}

// Opt the kernel in to more than 48 KB of dynamic shared memory.  Always the SAME ceiling (the
// largest request any launch of these kernels can make), never the launch's own size: two host
// threads launching models with different table sizes would otherwise lower the limit under
// each other between the attribute call and the launch.
constexpr size_t kSarSmemCeiling = kSarMaxStagedBytes * 4 / 3 + 16 * 1024;  // padded tables + result queues
template <class K>
cudaError_t allowSmem(K kernel, size_t bytes) {
  if (bytes > 48 * 1024)
    return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSarSmemCeiling);
  return cudaSuccess;
}

// getImageTime SAR:1113-1119, getSensorPosition(ImageCoord) SAR:1179-1184,
// getSensorVelocity(ImageCoord) SAR:1232-1238, per image point.
__global__ void sarSensorStateKernel(const __grid_constant__ SarState S, long n,
                                     const double *__restrict__ line, SensorStateOut o) {
  const SarTables T = stageSarTables<false>(S, nullptr);
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double time =
        o.atTime ? line[i] : S.startingEphemerisTime + (line[i] - 0.5) * S.exposureDuration;
    double pos[3], vel[3];
    sarPosVel(S, T, time, pos, vel);
    storeSensorState(o, i, time, pos, vel, 0u);
  }
}

}  // namespace

cudaError_t launchSarSensorState(const SarState &S, long n, const double *line,
                                 const SensorStateOut &o, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  const long blocks = (n + 255) / 256;
  sarSensorStateKernel<<<(int)(blocks < 148 * 8 ? blocks : 148 * 8), 256, 0, stream>>>(S, n, line, o);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchSarGroundToImage(const SarState &S, int smCount, long n, const double *x,
                                   const double *y, const double *z, double prec, double *line,
                                   double *samp, double *achieved, uint32_t *status,
                                   cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  size_t smemBytes;
  bool stage;
  // USGSCSM_B200_SAR_REFILL: lanes take a new point as soon as their search ends (1) or warps run
  // 32 searches in lockstep (0, default: see DESIGN.md); USGSCSM_B200_SAR_PAD: staged tables in rows
  // of 4 doubles; USGSCSM_B200_SAR_MINB: register budget as resident blocks per SM (2 or 3)
  static const int refill = sarEnv("USGSCSM_B200_SAR_REFILL", 0);
  static const int pad = sarEnv("USGSCSM_B200_SAR_PAD", 0);
  // USGSCSM_B200_SAR_INTERLEAVE=0: positions and velocities as two tables (rows of 3 doubles)
  static const int interleave = sarEnv("USGSCSM_B200_SAR_INTERLEAVE", 1);
  const int layout = pad ? kSarPadded : (interleave && sarInterleavable(S) ? kSarInterleaved : kSarPlain);
  static const int minb = sarEnv("USGSCSM_B200_SAR_MINB", 3);
  const size_t queueBytes = refill ? (kThreads / 32) * sizeof(SarWarpQueue) : 0;
  LaunchShape sh = sarShape(S, n, smCount, smemBytes, stage, queueBytes, minb >= 3 ? 3 : 2, layout);
  WorkCounter wc;
  cudaError_t e = cudaSuccess;
#define UB_SAR_G2I(KERNEL, STAGE, PAD, MINB)                                                      \
  do {                                                                                            \
    e = allowSmem(KERNEL<STAGE, PAD, MINB>, smemBytes);                                           \
    if (e == cudaSuccess) e = wc.acquire(stream, refill ? -1 : n);                                                 \
    if (e == cudaSuccess)                                                                         \
      KERNEL<STAGE, PAD, MINB><<<sh.blocks, sh.threads, smemBytes, stream>>>(                     \
          S, n, x, y, z, prec, line, samp, achieved, status, wc.ptr, grabFor(n));                 \
  } while (0)
#define UB_SAR_G2I_M(KERNEL, STAGE, PAD)            \
  do {                                              \
    if (minb >= 3) UB_SAR_G2I(KERNEL, STAGE, PAD, 3); \
    else UB_SAR_G2I(KERNEL, STAGE, PAD, 2);         \
  } while (0)
#define UB_SAR_G2I_K(KERNEL)                                                        \
  do {                                                                              \
    if (stage && layout == kSarPadded) UB_SAR_G2I_M(KERNEL, true, kSarPadded);      \
    else if (stage && layout == kSarInterleaved) UB_SAR_G2I_M(KERNEL, true, kSarInterleaved); \
    else if (stage) UB_SAR_G2I_M(KERNEL, true, kSarPlain);                          \
    else UB_SAR_G2I_M(KERNEL, false, kSarPlain);                                    \
  } while (0)
  if (refill)
    UB_SAR_G2I_K(sarGroundToImageRefillKernel);
  else
    UB_SAR_G2I_K(sarGroundToImageKernel);
#undef UB_SAR_G2I_K
#undef UB_SAR_G2I_M
#undef UB_SAR_G2I
  if (e != cudaSuccess) return e;
  countLaunch();
  return wc.release();
}

cudaError_t launchSarImageToGround(const SarState &S, int smCount, long n, const double *line,
                                   const double *samp, const GridSpec *grid, const double *height,
                                   int heightStride, double prec, double *x, double *y, double *z,
                                   double *achieved, uint32_t *status, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  size_t smemBytes;
  bool stage;
  LaunchShape sh = sarShape(S, n, smCount, smemBytes, stage);
  GridSpec g = grid ? *grid : GridSpec{0, 0, 0, 0, 1, 0};
  if (stage && smemBytes > 48 * 1024) {
    cudaError_t ea = grid ? allowSmem(sarImageToGroundKernel<true, true>, smemBytes)
                          : allowSmem(sarImageToGroundKernel<true, false>, smemBytes);
    if (ea != cudaSuccess) return ea;
  }
  WorkCounter wc;
  cudaError_t e0 = wc.acquire(stream, n);
  if (e0 != cudaSuccess) return e0;
#define UB_LAUNCH_SAR(STAGE, GRID)                                                          \
  do {                                                                                      \
    sarImageToGroundKernel<STAGE, GRID><<<sh.blocks, sh.threads, smemBytes, stream>>>(      \
        S, n, line, samp, g, height, heightStride, prec, x, y, z, achieved, status, wc.ptr, \
        grabFor(n));                                                                        \
  } while (0)
  if (stage && grid)
    UB_LAUNCH_SAR(true, true);
  else if (stage)
    UB_LAUNCH_SAR(true, false);
  else if (grid)
    UB_LAUNCH_SAR(false, true);
  else
    UB_LAUNCH_SAR(false, false);
#undef UB_LAUNCH_SAR
  countLaunch();
  return wc.release();
}

cudaError_t launchSarRemoteLocus(const SarState &S, int smCount, long n, const double *line,
                                 const double *samp, double prec, const LocusOut &o,
                                 cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  size_t smemBytes;
  bool stage;
  LaunchShape sh = sarShape(S, n, smCount, smemBytes, stage);
  if (stage) {
    cudaError_t e = allowSmem(sarRemoteLocusKernel<true>, smemBytes);
    if (e != cudaSuccess) return e;
    sarRemoteLocusKernel<true><<<sh.blocks, sh.threads, smemBytes, stream>>>(S, n, line, samp, prec, o);
  } else {
    sarRemoteLocusKernel<false><<<sh.blocks, sh.threads, 0, stream>>>(S, n, line, samp, prec, o);
  }
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchGroundPartialsSar(const SarState &S, int smCount, long n, const double *x,
                                    const double *y, const double *z, double *partials,
                                    uint32_t *status, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  size_t smemBytes;
  bool stage;
  LaunchShape sh = sarShape(S, n, smCount, smemBytes, stage);
  if (stage) {
    cudaError_t e = allowSmem(sarGroundPartialsKernel<true>, smemBytes);
    if (e != cudaSuccess) return e;
  }
  WorkCounter wc;
  cudaError_t e = wc.acquire(stream, n);
  if (e != cudaSuccess) return e;
  if (stage)
    sarGroundPartialsKernel<true><<<sh.blocks, sh.threads, smemBytes, stream>>>(
        S, n, x, y, z, partials, status, wc.ptr, grabFor(n));
  else
    sarGroundPartialsKernel<false><<<sh.blocks, sh.threads, 0, stream>>>(
        S, n, x, y, z, partials, status, wc.ptr, grabFor(n));
  countLaunch();
  return wc.release();
}

cudaError_t launchSarProximateLocus(const SarState &S, int smCount, long n, const double *line,
                                    const double *samp, const double *x, const double *y,
                                    const double *z, const LocusOut &o, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  size_t smemBytes;
  bool stage;
  LaunchShape sh = sarShape(S, n, smCount, smemBytes, stage);
  if (stage) {
    cudaError_t e = allowSmem(sarProximateLocusKernel<true>, smemBytes);
    if (e != cudaSuccess) return e;
    sarProximateLocusKernel<true><<<sh.blocks, sh.threads, smemBytes, stream>>>(S, n, line, samp, x, y, z, o);
  } else {
    sarProximateLocusKernel<false><<<sh.blocks, sh.threads, 0, stream>>>(S, n, line, samp, x, y, z, o);
  }
  countLaunch();
  return cudaGetLastError();
}

}  // namespace ub200
