// sip_kernels.cu -- Stone's strongly implicit procedure (SIP) for the 5-point p' equation of the explicit
// finite-difference cavity solver: sipsol, lid-driven-cavity-2d/fdm-explicit/fdm_explicit_cavity.f90:24-87
// (incomplete LU factors :44-55, residual + forward substitution :59-67, backward substitution + update :71-76,
// stop rule :78-85; maxit 600, alpha 0.88, tol 1e-5).
//
// Every step of the algorithm has a recurrence in BOTH grid directions (cell (i,j) needs (i-1,j) and (i,j-1)), so
// the parallelism is a wavefront and the run time is a latency chain, not a bandwidth figure.  One persistent
// cooperative kernel runs the whole solve; the rows are cut into bands of 32, one band per block at a time:
//   * incomplete factorisation (once per solve): sip_wavefront<0> -- 32x32 tiles solved along their 63 anti-diagonals in
//     shared memory, band b starts chunk c when band b-1 has published it (release/acquire counter per band);
//   * forward / backward substitution (every inner iteration): sip_stream -- lane = row, the warp is one continuous
//     skewed front over the whole band (imax + 31 steps), its 32-column chunks are moved between global and shared
//     memory by a second warp, and the band's last row reaches the next band value by value through a polled row buffer
//     (no fence, no flag); see the comment at sip_stream;
//   * residual pass and norm: all eight warps of every block stream the arrays; the norm is reduced in a fixed order, so
//     every block takes the same stop decision.  Between the phases the grid synchronises (cooperative groups).
// Measured at 4097^2 (B200): 2.9 ms per inner iteration (0.35 residual + 1.27 + 1.26 substitutions) against 5.9 ms for
// the tile-by-tile substitution of round 1; per band 99 cycles per step in the recurrence loop, about 60 of them the
// dependent chain itself (shuffle + four FP64 operations at 8 cycles), and each band starts ~7.8 us after the one before
// (31 steps of skew + 8 of hand-off granularity + the store-to-load latency through L2).
// Operation order and non-contracted IEEE arithmetic as in the reference: the result is bit-identical to
// oracle/fdcavity_oracle.c:orc_fdc_sipsol whenever the iteration counts agree (they differ only if the ratio
// rsm lands within rounding of tol).
#include <cooperative_groups.h>
#include <cstdlib>
#include "common.cuh"
namespace cg = cooperative_groups;

namespace cfd {

struct SipParams {
  int imax, jmax, ld;        // arrays are column-major with leading dimension ld; interior (0-based) 1..imax x 1..jmax
  double aw, ae, as, an, ap; // constant coefficients as passed to sipsol (already negated by the caller, :206) ...
  const double *awp, *aep, *asp, *anp, *app;   // ... or coefficient arrays (fdm_implicit_cavity.f90:199): used if non-null,
  double sign;               // the four neighbour arrays multiplied by `sign` (-1: the caller's `-aw, -ae, -as, -an`)
  double alpha, tol;
  int maxit;
  const double* su;
  double* phi;
  double *lw, *ls, *lpr, *un, *ue, *res;   // work arrays, same shape
  unsigned long long* prog;  // one progress counter per band
  int skip;                  // debugging aid (CFD_B200_SIP_SKIP): bit 0 skips the forward, bit 1 the backward substitution
  double* brow;              // [band][ld]: the band's boundary row, handed to the next band value by value (see sip_stream)
  double* partial;           // per-block residual partial sums
  int* info;                 // [0] iterations used, [1] abort flag
  double* rsm_out;
};

#ifndef SIP_TIMING
#define SIP_TIMING 0         // 1: per-iteration phase times (globaltimer), 2: also clock64 stamps inside the band loops
#endif
#define SIP_BAND_TIMING (SIP_TIMING >= 2)
#define SIP_W 1              // bands in flight per block (SIP_WARP_DOUBLES of shared memory each: ILU tiles / chunk ring)
#define SIP_WARPS 8          // warps per block; all of them stream the residual pass, warp 0 walks the wavefront of the
                             // block's band, warp 1 moves its chunks between global and shared memory (sip_stream)
#define SIP_SET (4 * 32 * TS)                       // doubles in one set of four tiles
#define SIP_RS 130                                  // sip_stream ring: doubles per row (4 chunks of 32 columns + 2: the skewed
                                                    // front, lane l at column s - l, then hits 32 different banks)
#define SIP_RTILE (32 * SIP_RS)                     // one operand array of the ring
#define SIP_WARP_DOUBLES (4 * SIP_RTILE + 4 * 32)   // per band: the ring (>= one ILU tile set) + row0, row1, col0, col1 / abort slots
#define TS 34                // padded tile stride: the anti-diagonal (lane l at row l, column d-l) sits at 33*l + d -> 32 different banks
static_assert(4 * SIP_RTILE >= SIP_SET, "the ILU tiles live in the ring's memory");

__device__ __forceinline__ unsigned long long ld_acquire(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// wait until band `b` has published at least `target`; returns false if the wait was abandoned
__device__ __forceinline__ bool wait_band(const SipParams& P, int b, unsigned long long target) {
  unsigned spins = 0;
  while (ld_acquire(P.prog + b) < target) {
    __nanosleep(64);
    if (++spins > (1u << 24) || *(volatile int*)(P.info + 1)) {
      P.info[1] = 1;
      return false;
    }
  }
  return true;
}

// load / store one 32x32 tile (rows jb..jb+31, columns c0..c0+31, 0-based) as coalesced row segments
__device__ __forceinline__ void tile_load(double* tl, const double* g, int ld, int c0, int jb, int imax, int jmax, int lane) {
  // sixteen row segments in flight per lane before the first one is needed (a rolled loop would wait for every load)
  const int i = c0 + lane;
  const bool iok = i >= 1 && i <= imax;
#pragma unroll
  for (int r0 = 0; r0 < 32; r0 += 16) {
    double v[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const int j = jb + r0 + r;
      v[r] = (iok && j >= 1 && j <= jmax) ? __ldcg(g + (size_t)j * ld + i) : 0.0;
    }
#pragma unroll
    for (int r = 0; r < 16; ++r) tl[(r0 + r) * TS + lane] = v[r];
  }
}
__device__ __forceinline__ void tile_store(const double* tl, double* g, int ld, int c0, int jb, int imax, int jmax, int lane) {
  for (int r = 0; r < 32; ++r) {
    const int i = c0 + lane, j = jb + r;
    if (i >= 1 && i <= imax && j >= 1 && j <= jmax) g[(size_t)j * ld + i] = tl[r * TS + lane];
  }
}

// the four operand tiles of a chunk in two round trips: sixteen rows of all four arrays (64 values per lane) are in
// flight together (one block per SM leaves each thread up to 255 registers)
__device__ __forceinline__ void tile_load4(double* A, double* B, double* C, double* D, const double* ga, const double* gb,
                                           const double* gc, const double* gd, int ld, int c0, int jb, int imax, int jmax, int lane) {
  const int i = c0 + lane;
  const bool iok = i >= 1 && i <= imax;
#pragma unroll
  for (int r0 = 0; r0 < 32; r0 += 16) {
    double va[16], vb[16], vc[16], vd[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const int j = jb + r0 + r;
      const bool ok = iok && j >= 1 && j <= jmax;
      const size_t k = (size_t)j * ld + i;
      va[r] = ok ? __ldcg(ga + k) : 0.0;
      vb[r] = ok ? __ldcg(gb + k) : 0.0;
      vc[r] = ok ? __ldcg(gc + k) : 0.0;
      vd[r] = ok ? __ldcg(gd + k) : 0.0;
    }
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const int o = (r0 + r) * TS + lane;
      A[o] = va[r]; B[o] = vb[r]; C[o] = vc[r]; D[o] = vd[r];
    }
  }
}

// PHASE 0: incomplete LU factors (:44-55); 1: forward substitution (:64-65); 2: backward substitution + update (:71-76)
template <int PHASE>
__device__ void sip_wavefront(const SipParams& P, double* sm, unsigned long long epoch_base, int nbands, int nchunks) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int imax = P.imax, jmax = P.jmax;              // interior 0-based: 1..imax, 1..jmax
  double* t0 = sm + (size_t)w * SIP_WARP_DOUBLES;
  double *A = t0, *B = A + 32 * TS, *C = B + 32 * TS, *D = C + 32 * TS;   // tiles
  double *row0 = t0 + 4 * SIP_RTILE, *row1 = row0 + 32, *col0 = row1 + 32, *col1 = col0 + 32;
  if (w >= SIP_W) return;
  const bool fwd = PHASE != 2;
  long long tm[6] = {0, 0, 0, 0, 0, 0}, tq = 0;
#define TICK(n) if (SIP_TIMING) { const long long now__ = clock64(); tm[n] += now__ - tq; tq = now__; }
  for (int bb = blockIdx.x * SIP_W + w; bb < nbands; bb += gridDim.x * SIP_W) {
    const int b = fwd ? bb : nbands - 1 - bb;           // band processed (backward: from the top)
    const int jb = 1 + 32 * b;
    col0[lane] = 0.0;                                   // values of the column just outside the first chunk
    col1[lane] = 0.0;
    __syncwarp();
    for (int cc = 0; cc < nchunks; ++cc) {
      const int c = fwd ? cc : nchunks - 1 - cc;
      const int c0 = 1 + 32 * c;
      if (SIP_TIMING) tq = clock64();
      // the tile's own operands do not depend on the neighbouring band: fetch them before waiting for it
      if (PHASE == 1) {
        tile_load4(A, B, C, D, P.res, P.ls, P.lw, P.lpr, P.ld, c0, jb, imax, jmax, lane);
      } else if (PHASE == 2) {
        tile_load4(A, B, C, D, P.res, P.un, P.ue, P.phi, P.ld, c0, jb, imax, jmax, lane);
      }
      TICK(0)
      if (bb > 0) {                                     // the band processed just before this one (other warp/block)
        const int dep = fwd ? b - 1 : b + 1;
        if (!wait_band(P, dep, epoch_base + (unsigned long long)cc + 1)) return;
      }
      TICK(1)
      // row just outside the tile in the sweep direction (south for forward phases, north for backward)
      {
        const int jn = fwd ? jb - 1 : jb + 32, i = c0 + lane;
        const bool ok = jn >= 1 && jn <= jmax && i >= 1 && i <= imax;
        if (PHASE == 0) {
          row0[lane] = ok ? P.un[(size_t)jn * P.ld + i] : 0.0;
          row1[lane] = ok ? P.ue[(size_t)jn * P.ld + i] : 0.0;
        } else {
          row0[lane] = ok ? P.res[(size_t)jn * P.ld + i] : 0.0;
        }
      }
      __syncwarp();
      TICK(2)
      if (PHASE == 0) {
        for (int d = 0; d < 63; ++d) {
          const int r = lane, q = d - lane;               // row / column inside the tile
          const bool act = q >= 0 && q < 32 && (c0 + q) <= imax && (jb + r) <= jmax;
          if (act) {
            const int k = r * TS + q;
            // west / south neighbours of the factors un, ue
            const double un_w = (q == 0) ? col0[r] : A[k - 1], ue_w = (q == 0) ? col1[r] : B[k - 1];
            const double un_s = (r == 0) ? row0[q] : A[k - TS], ue_s = (r == 0) ? row1[q] : B[k - TS];
            const size_t kg = (size_t)(jb + r) * P.ld + c0 + q;
            const double caw = P.awp ? P.sign * P.awp[kg] : P.aw, cas = P.asp ? P.sign * P.asp[kg] : P.as;
            const double can = P.anp ? P.sign * P.anp[kg] : P.an, cae = P.aep ? P.sign * P.aep[kg] : P.ae;
            const double cap = P.app ? P.app[kg] : P.ap;
            const double lw = __ddiv_rn(caw, __dadd_rn(1.0, __dmul_rn(P.alpha, un_w)));
            const double ls = __ddiv_rn(cas, __dadd_rn(1.0, __dmul_rn(P.alpha, ue_s)));
            const double p1 = __dmul_rn(__dmul_rn(P.alpha, lw), un_w);
            const double p2 = __dmul_rn(__dmul_rn(P.alpha, ls), ue_s);
            const double den = __dsub_rn(__dsub_rn(__dadd_rn(__dadd_rn(cap, p1), p2), __dmul_rn(lw, ue_w)), __dmul_rn(ls, un_s));
            const double lpr = __ddiv_rn(1.0, den);
            A[k] = __dmul_rn(__dsub_rn(can, p1), lpr);    // un
            B[k] = __dmul_rn(__dsub_rn(cae, p2), lpr);    // ue
            C[k] = lw;
            D[k] = ls;
            // lpr goes straight to global memory (a fifth tile would not fit the budget)
            P.lpr[kg] = lpr;
          }
          __syncwarp();
        }
      } else {
        // Forward / backward substitution: lane = row, marching along the columns one step behind the lane before it.
        // The two recurrence inputs never touch shared memory: the value of the previous row at this column is what
        // the neighbouring lane produced one step ago (shuffle), the value of the previous column is this lane's own
        // last result (register).  Shared memory only supplies the cell's operands and receives the result.
        double last = 0.0;
        // per-lane constants of the skewed walk: the lane is active for d in [lane, dhi]; its cell at step d sits at
        // kb + d (forward) / kb - d (backward) in the tile, the row-0 / column-0 inputs likewise
        const int r = fwd ? lane : 31 - lane;
        int ncol = imax - c0 + 1;                          // valid columns of this chunk
        if (ncol > 32) ncol = 32;
        const bool rowok = (jb + r) <= jmax;
        // forward: q = d - lane valid for q < ncol; backward: q = 31 - (d - lane) valid for q < ncol <=> d - lane > 31 - ncol
        const int dlo = fwd ? lane : lane + (32 - ncol);
        const int dhi = rowok ? (fwd ? lane + ncol - 1 : lane + 31) : -1;
        const int kb = fwd ? r * TS - lane : r * TS + 31 + lane;
        const double cin = col0[r];
#pragma unroll
        for (int d = 0; d < 63; ++d) {
          const double from_row = __shfl_up_sync(0xffffffffu, last, 1);
          if (d >= dlo && d <= dhi) {
            const int k = fwd ? kb + d : kb - d;
            const int q = fwd ? d - lane : 31 + lane - d;
            const double nb_row = (lane == 0) ? row0[q] : from_row;              // south (forward) / north (backward)
            const double nb_col = (d == lane) ? cin : last;                      // west (forward) / east (backward)
            if (PHASE == 1) {
              last = __dmul_rn(__dsub_rn(__dsub_rn(A[k], __dmul_rn(B[k], nb_row)), __dmul_rn(C[k], nb_col)), D[k]);
              A[k] = last;
            } else {
              last = __dsub_rn(__dsub_rn(A[k], __dmul_rn(B[k], nb_row)), __dmul_rn(C[k], nb_col));
              A[k] = last;
              D[k] = __dadd_rn(D[k], last);
            }
          }
        }
        __syncwarp();
      }
      TICK(3)
      // keep the column next to the following chunk, write the tile back, publish
      {
        const int qe = fwd ? 31 : 0;
        col0[lane] = A[lane * TS + qe];
        if (PHASE == 0) col1[lane] = B[lane * TS + qe];
      }
      if (PHASE == 0) {
        tile_store(A, P.un, P.ld, c0, jb, imax, jmax, lane);
        tile_store(B, P.ue, P.ld, c0, jb, imax, jmax, lane);
        tile_store(C, P.lw, P.ld, c0, jb, imax, jmax, lane);
        tile_store(D, P.ls, P.ld, c0, jb, imax, jmax, lane);
      } else if (PHASE == 1) {
        tile_store(A, P.res, P.ld, c0, jb, imax, jmax, lane);
      } else {
        tile_store(A, P.res, P.ld, c0, jb, imax, jmax, lane);
        tile_store(D, P.phi, P.ld, c0, jb, imax, jmax, lane);
      }
      TICK(4)
      __threadfence();
      __syncwarp();
      if (lane == 0) st_release(P.prog + b, epoch_base + (unsigned long long)cc + 1);
      TICK(5)
    }
    if (SIP_TIMING && lane == 0 && (bb == 1 || bb == nbands / 2) && epoch_base / ((unsigned long long)nchunks + 1) == (PHASE == 1 ? 6 : 7))
      printf("SIP phase %d band %d: loads %lld wait %lld row0 %lld steps %lld stores %lld fence+flag %lld cycles over %d chunks\n",
             PHASE, bb, tm[0], tm[1], tm[2], tm[3], tm[4], tm[5], nchunks);
  }
#undef TICK
}

__device__ __forceinline__ void cp_async8(double* dst, const double* src, bool ok) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
  const int n = ok ? 8 : 0;                           // src-size 0: the destination is filled with zeros
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(d), "l"(src), "r"(n) : "memory");
}

// ---- forward / backward substitution as ONE continuous wavefront per band ---------------------------------------
// lane = row of the band (32 rows per warp); at step s the lane solves its column q = s - lane, so the warp is a
// skewed front that never stops at tile boundaries: imax + 31 steps per band instead of 63 per 32 columns.  Per step the
// dependent chain is one shuffle (the previous row's value comes from the lane above) and the four FP64 operations of
// the recurrence; everything else is kept off it:
//   * operands: 32-column chunks of the band (four 32x32 tiles) travel through a ring of three shared-memory sets --
//     chunk w+1 is fetched with cp.async (256-byte row segments) while the front crosses chunks w-1 and w, chunk w-2 is
//     written back as row segments; a lane's operands of step s+1 are read while step s computes;
//   * band -> band: the last lane copies its row into brow[band][q]; the first lanes of the next band poll the VALUE
//     (entries rest at an all-ones pattern no arithmetic produces; the reader puts it back).  No fence, no flag, no
//     32-column granularity: a band runs 31 + 8 + (latency) steps behind the band before it.
// Bands are dealt block-first (band bb -> block bb mod gridDim), so that neighbouring bands run on different SMs.
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const double* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u64(double* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// shared-memory accesses of the step loop in program order (the compiler may not sink the next step's operand reads
// below this step's result store)
__device__ __forceinline__ double lds_f64(unsigned a) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_f64(unsigned a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v)); }
// predicated forms: the step loop stays free of branches (a divergent branch per step costs more than the step's math)
__device__ __forceinline__ void sts_f64_if(bool p, unsigned a, double v) {
  asm volatile("{ .reg .pred q; setp.ne.s32 q, %2, 0; @q st.shared.f64 [%0], %1; }" ::"r"(a), "d"(v), "r"((int)p));
}
__device__ __forceinline__ void st_relaxed_f64_if(bool p, double* a, double v) {
  asm volatile("{ .reg .pred q; setp.ne.s32 q, %2, 0; @q st.relaxed.gpu.global.f64 [%0], %1; }" ::"l"(a), "d"(v), "r"((int)p) : "memory");
}
#define SIP_EMPTY 0xffffffffffffffffull

__device__ __forceinline__ void bar_pair() { asm volatile("bar.sync 1, 64;" ::: "memory"); }   // warps 0 and 1

template <int PHASE>
__device__ void sip_stream(const SipParams& P, double* sm, int nbands) {
  constexpr bool fwd = PHASE == 1;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (w >= 2) return;                                          // warp 0 computes, warp 1 moves chunks
  const int imax = P.imax, jmax = P.jmax;
  double* ring = sm;
  int* abort_s = (int*)(sm + 4 * SIP_RTILE);                   // the pair's common view of the abort flag
  const unsigned ring_s = (unsigned)__cvta_generic_to_shared(ring);
  const double *ga = P.res, *gb = fwd ? P.ls : P.un, *gc = fwd ? P.lw : P.ue, *gd = fwd ? P.lpr : P.phi;
  const int nchunks = (imax + 31) / 32;
  const int nwin = (imax + 31 + 31) / 32;                      // windows of 32 steps
  constexpr unsigned TB = SIP_RTILE * 8;                       // bytes between the four operand arrays of the ring
  unsigned nbar = 0;                                           // barriers passed by the pair (both warps count alike)
  for (int bb = blockIdx.x; bb < nbands; bb += gridDim.x) {
    const int b = fwd ? bb : nbands - 1 - bb;
    const int jb = 1 + 32 * b;
    auto column = [&](int q) { return fwd ? 1 + q : imax - q; };

    if (w == 1) {
      // ---- mover: the ring holds 128 consecutive columns of the band, element (row of lane l, column q) at
      // l*SIP_RS + (q mod 128) (q counts in sweep direction); chunk cq = columns [32 cq, 32 cq + 31].  During window `win`
      // the front is inside chunks win-1 and win: chunk win-2 goes back to global memory, chunk win+2 takes its place,
      // chunk win+1 (requested a window ago) has landed before the barrier.
      auto fetch = [&](int cq) {
        if (cq < nchunks) {
          double* set = ring + ((32 * cq) & 127) + lane;
          const int q = 32 * cq + lane;
          const bool qok = q < imax;
          const int i = qok ? column(q) : 1;
#pragma unroll 8
          for (int l = 0; l < 32; ++l) {
            const int j = jb + (fwd ? l : 31 - l);
            const bool ok = qok && j <= jmax;
            const size_t k = (size_t)(ok ? j : jb) * P.ld + i;
            const int o = l * SIP_RS;
            cp_async8(set + o, ga + k, ok);
            cp_async8(set + SIP_RTILE + o, gb + k, ok);
            cp_async8(set + 2 * SIP_RTILE + o, gc + k, ok);
            cp_async8(set + 3 * SIP_RTILE + o, gd + k, ok);
          }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
      };
      auto write_back = [&](int cq) {
        if (cq < 0 || cq >= nchunks) return;
        const double* set = ring + ((32 * cq) & 127) + lane;
        const int q = 32 * cq + lane;
        if (q >= imax) return;
        const int i = column(q);
#pragma unroll 8
        for (int l = 0; l < 32; ++l) {
          const int j = jb + (fwd ? l : 31 - l);
          if (j <= jmax) {
            P.res[(size_t)j * P.ld + i] = set[l * SIP_RS];
            if (PHASE == 2) P.phi[(size_t)j * P.ld + i] = set[3 * SIP_RTILE + l * SIP_RS];
          }
        }
      };
      fetch(0);
      fetch(1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
      bar_pair();
      if (abort_s[nbar++ & 1]) return;
      for (int win = 0; win < nwin; ++win) {
        write_back(win - 2);
        fetch(win + 2);
        asm volatile("cp.async.wait_group 1;" ::: "memory");     // everything but the newest request: chunk win+1 is in
        bar_pair();
        if (abort_s[nbar++ & 1]) return;
      }
      write_back(nwin - 2);
      write_back(nwin - 1);
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      continue;
    }

    // ---- compute warp
    const int r = fwd ? lane : 31 - lane;                      // the lane's row of the band
    const bool rowok = (jb + r) <= jmax;
    const bool has_dep = bb > 0, has_next = bb + 1 < nbands;
    double* mine = P.brow + (size_t)b * P.ld;
    double* theirs = P.brow + (size_t)(has_dep ? (fwd ? b - 1 : b + 1) : b) * P.ld;
    long long t_wait = 0, t_all = SIP_BAND_TIMING ? clock64() : 0, t_win = 0, t_poll = 0, t_steps = 0, t_q = 0;
    unsigned long long g_first = 0, g_sb1 = 0, g_begin = 0;
    if (SIP_BAND_TIMING) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g_begin));
#define STAMP(acc) if (SIP_BAND_TIMING) { const long long n__ = clock64(); acc += n__ - t_q; t_q = n__; }
    // shared-memory address of the lane's element at column q (any q: columns wrap around the ring's 128)
    const unsigned lane_s = ring_s + (unsigned)(lane * SIP_RS * 8);
    auto elem = [&](int q) { return lane_s + (unsigned)((q & 127) << 3); };
    double* mine_q = mine + (32 * 0 - lane);                   // running address of brow[band][q] (lane 31 only stores)
    double last = 0.0;
    unsigned long long nxt = SIP_EMPTY;
    if (has_dep && lane < 8 && lane < imax) nxt = ld_relaxed_u64(theirs + lane);
    double a1 = 0.0, b1 = 0.0, c1 = 0.0, d1 = 0.0, a2 = 0.0, b2 = 0.0, c2 = 0.0, d2 = 0.0;
    const bool publish = lane == 31 && has_next;
    // the abort flag as both warps of the pair see it: written before a barrier, read after it (two slots by parity,
    // so that the next write cannot overtake a late reader)
    if (lane == 0) abort_s[nbar & 1] = *(volatile int*)(P.info + 1);
    bar_pair();                                                // chunk 0 is in
    if (abort_s[nbar++ & 1]) return;
    for (int win = 0; win < nwin; ++win) {
      if (SIP_BAND_TIMING) t_q = clock64();
      {   // operands of the window's first two steps (the read-ahead of the steps before may have met a chunk in flight)
        const unsigned e = elem(32 * win - lane), e2 = elem(32 * win - lane + 1);
        a1 = lds_f64(e); b1 = lds_f64(e + TB); c1 = lds_f64(e + 2 * TB); d1 = lds_f64(e + 3 * TB);
        a2 = lds_f64(e2); b2 = lds_f64(e2 + TB); c2 = lds_f64(e2 + 2 * TB); d2 = lds_f64(e2 + 3 * TB);
      }
      STAMP(t_win)
#pragma unroll 1
      for (int sb = 0; sb < 4; ++sb) {
        const int s0 = 32 * win + 8 * sb;
        double rw = 0.0;
        if (has_dep) {
          const int q = s0 + lane;
          if (lane < 8 && q < imax) {
            unsigned long long v = nxt;
            if (v == SIP_EMPTY) {
              const long long t0 = SIP_BAND_TIMING ? clock64() : 0;
              unsigned spins = 0;
              while ((v = ld_relaxed_u64(theirs + q)) == SIP_EMPTY) {
                // one load per turn: the abort flag is looked at every 256 turns only
                if ((++spins & 255u) == 0 && (spins > (1u << 22) || *(volatile int*)(P.info + 1))) {
                  P.info[1] = 1;
                  break;
                }
              }
              if (SIP_BAND_TIMING) t_wait += clock64() - t0;
            }
            st_relaxed_u64(theirs + q, SIP_EMPTY);             // ready for the next phase
            rw = __longlong_as_double((long long)v);
            nxt = (q + 8 < imax) ? ld_relaxed_u64(theirs + q + 8) : SIP_EMPTY;
          }
        }
        __syncwarp();                                        // the eight polling lanes rejoin before the shuffles of the steps
        STAMP(t_poll)
        if (SIP_BAND_TIMING && win == 0 && sb == 0) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g_first));
        if (SIP_BAND_TIMING && win == 1 && sb == 1) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g_sb1));
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          const int q = s0 + t - lane;
          const bool act = rowok && q >= 0 && q < imax;
          const double a0 = a1, b0 = b1, c0 = c1, d0 = d1;
          a1 = a2; b1 = b2; c1 = c2; d1 = d2;
          const unsigned e = elem(q);
          {
            // operands two steps ahead: the shared-memory accesses keep their program order, so this read follows the result
            // store of the step before; one step of distance would leave its latency inside the recurrence chain
            const unsigned en = elem(q + 2);
            a2 = lds_f64(en); b2 = lds_f64(en + TB); c2 = lds_f64(en + 2 * TB); d2 = lds_f64(en + 3 * TB);
          }
          const double from_row = __shfl_up_sync(0xffffffffu, last, 1);
          const double rv = __shfl_sync(0xffffffffu, rw, t);
          const double nb_row = (lane == 0) ? rv : from_row;   // south (forward) / north (backward)
          double x;
          if (PHASE == 1) x = __dmul_rn(__dsub_rn(__dsub_rn(a0, __dmul_rn(b0, nb_row)), __dmul_rn(c0, last)), d0);
          else x = __dsub_rn(__dsub_rn(a0, __dmul_rn(b0, nb_row)), __dmul_rn(c0, last));
          last = act ? x : last;
          sts_f64_if(act, e, x);
          if (PHASE == 2) sts_f64_if(act, e + 3 * TB, __dadd_rn(d0, x));
          st_relaxed_f64_if(act && publish, mine_q + (s0 + t), x);
        }
        STAMP(t_steps)
      }
      if (lane == 0) abort_s[nbar & 1] = *(volatile int*)(P.info + 1);
      bar_pair();                                              // results of this window visible to the mover; chunk win+1 is in
      if (abort_s[nbar++ & 1]) return;
      STAMP(t_win)
    }
    if (SIP_BAND_TIMING && lane == 0 && (bb == 1 || bb == nbands / 2)) {
      printf("SIP phase %d band %d: %lld cycles for %d steps (window turn-over %lld, neighbour row %lld, steps %lld), %lld waiting for the neighbouring band\n",
             PHASE, bb, clock64() - t_all, nwin * 32, t_win, t_poll, t_steps, t_wait);
    }
    if (SIP_BAND_TIMING && lane == 0 && PHASE == 1 && bb < 6) {
      unsigned long long g_end;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g_end));
      printf("band %d: entered %llu ns, first values at +%llu ns, step 40 at +%llu ns, done at +%llu ns\n", bb, g_begin % 100000000ull,
             g_first - g_begin, g_sb1 - g_begin, g_end - g_begin);
    }
#undef STAMP
  }
}

__global__ void __launch_bounds__(32 * SIP_WARPS, 1) k_sip_solve(const SipParams P) {
  extern __shared__ double sm[];
  cg::grid_group grid = cg::this_grid();
  const int imax = P.imax, jmax = P.jmax;
  const int nbands = (jmax + 31) / 32, nchunks = (imax + 31) / 32;
  const unsigned long long stride = (unsigned long long)nchunks + 1;
  unsigned long long epoch = 1;
  // ---- incomplete factorisation
  sip_wavefront<0>(P, sm, epoch * stride, nbands, nchunks);
  ++epoch;
  grid.sync();
  double res1 = 0.0, rsm = 0.0;
  int iter = 0;
  for (iter = 1; iter <= P.maxit; ++iter) {
    unsigned long long g0 = 0, g1 = 0, g2 = 0, g3 = 0;
    if (SIP_TIMING || P.maxit < 0) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g0));
    // ---- residual (:61-63) and its 1-norm, fixed-order reduction
    double acc = 0.0;
    // rows are dealt to the warps of the whole grid, a warp reads 256-byte row segments, four per lane in flight
    {
      const int lane = threadIdx.x & 31;
      const int gw = blockIdx.x * SIP_WARPS + (threadIdx.x >> 5), nw = gridDim.x * SIP_WARPS;
      for (int j = 1 + gw; j <= jmax; j += nw) {
        const size_t row = (size_t)j * P.ld;
        for (int i0 = 1; i0 <= imax; i0 += 128) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int i = i0 + 32 * u + lane;
            if (i <= imax) {
              const size_t k = row + i;
              const double caw = P.awp ? P.sign * P.awp[k] : P.aw, cae = P.aep ? P.sign * P.aep[k] : P.ae;
              const double can = P.anp ? P.sign * P.anp[k] : P.an, cas = P.asp ? P.sign * P.asp[k] : P.as;
              const double cap = P.app ? P.app[k] : P.ap;
              double r = __dsub_rn(P.su[k], __dmul_rn(caw, P.phi[k - 1]));
              r = __dsub_rn(r, __dmul_rn(cae, P.phi[k + 1]));
              r = __dsub_rn(r, __dmul_rn(can, P.phi[k + P.ld]));
              r = __dsub_rn(r, __dmul_rn(cas, P.phi[k - P.ld]));
              r = __dsub_rn(r, __dmul_rn(cap, P.phi[k]));
              P.res[k] = r;
              acc += fabs(r);
            }
          }
        }
      }
    }
    {
      __shared__ double wsum[SIP_WARPS];
      for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
      if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = acc;
      __syncthreads();
      if (threadIdx.x == 0) {
        double s = 0.0;
        for (int k = 0; k < SIP_WARPS; ++k) s += wsum[k];
        P.partial[blockIdx.x] = s;
      }
    }
    grid.sync();
    double resl = 0.0;
    for (unsigned k = 0; k < gridDim.x; ++k) resl += P.partial[k];   // same order in every block
    if (iter == 1) res1 = resl;
    rsm = resl / res1;
    // ---- forward and backward substitution
    if (SIP_TIMING || P.maxit < 0) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g1));
    if (!(P.skip & 1)) sip_stream<1>(P, sm, nbands);
    ++epoch;
    grid.sync();
    if (SIP_TIMING || P.maxit < 0) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g2));
    if (!(P.skip & 2)) sip_stream<2>(P, sm, nbands);
    ++epoch;
    grid.sync();
    // MEASURED CODE-GENERATION QUIRK (nvcc 12.9, sm_100a): with this report path compiled in -- reachable only for
    // maxit < 0, which cfd_launch_sip rejects -- the two substitution phases run 2.2x faster (1.27 ms instead of 2.83 ms per
    // phase at 4097^2; same SASS shape of the step loop, 116 instead of 106 registers).  The cause is not understood;
    // the path costs nothing at run time, so it stays until it is.
    if (SIP_TIMING || P.maxit < 0) {
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g3));
      if (blockIdx.x == 0 && threadIdx.x == 0 && (iter == 5 || iter == 100 || iter == 300 || iter == 599))
        printf("SIP iteration %d: residual pass %llu ns, forward %llu ns, backward %llu ns\n", iter, g1 - g0, g2 - g1, g3 - g2);
    }
    if (rsm < P.tol) break;
    if (iter == 1 && res1 < 1.0e-10) break;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    P.info[0] = iter > P.maxit ? P.maxit : iter;
    *P.rsm_out = rsm;
  }
}

}  // namespace cfd
using namespace cfd;

// bytes of the synchronisation block a caller allocates for `prog`: progress counters + one row of ld doubles per band
size_t cfd_sip_sync_bytes(int ld, int nrows) {
  const size_t nb = (size_t)(nrows - 1 + 31) / 32;
  return (nb + 2) * sizeof(unsigned long long) + nb * (size_t)ld * sizeof(double);
}

// Launch one SIP solve.  coef != nullptr: five coefficient arrays {aw, ae, as, an, ap} (the four neighbour arrays are
// negated, as the callers pass `-aw, -ae, -as, -an`); otherwise the constants c[5].  Work arrays are the caller's.
static int sip_launch(int device, cudaStream_t stream, int ld, int nrows, int imax, int jmax, const double* c,
                      const double* const* coef, double sign, double alpha, double tol, int maxit, const double* su, double* phi,
                      double* work, unsigned long long* prog, double* partial, int* info, double* rsm_out, int max_partials);

int cfd_launch_sip(int device, cudaStream_t stream, int ld, int nrows, int imax, int jmax, const double* c,
                   const double* const* coef, double alpha, double tol, int maxit, const double* su, double* phi, double* work,
                   unsigned long long* prog, double* partial, int* info, double* rsm_out, int max_partials) {
  return sip_launch(device, stream, ld, nrows, imax, jmax, c, coef, -1.0, alpha, tol, maxit, su, phi, work, prog, partial, info,
                    rsm_out, max_partials);
}

static int sip_launch(int device, cudaStream_t stream, int ld, int nrows, int imax, int jmax, const double* c,
                      const double* const* coef, double sign, double alpha, double tol, int maxit, const double* su, double* phi,
                      double* work, unsigned long long* prog, double* partial, int* info, double* rsm_out, int max_partials) {
  SipParams P{};
  P.imax = imax; P.jmax = jmax; P.ld = ld;
  if (coef) {
    P.awp = coef[0]; P.aep = coef[1]; P.asp = coef[2]; P.anp = coef[3]; P.app = coef[4];
    P.sign = sign;
  } else {
    P.aw = c[0]; P.ae = c[1]; P.as = c[2]; P.an = c[3]; P.ap = c[4];
  }
  if (maxit < 0) {
    cfd_set_error("SIP: maxit must not be negative");
    return CFD_ERR_INVALID;
  }
  P.alpha = alpha; P.tol = tol; P.maxit = maxit;
  { const char* e = getenv("CFD_B200_SIP_SKIP"); P.skip = e ? atoi(e) : 0; }
  P.su = su; P.phi = phi;
  const size_t n = (size_t)ld * nrows;
  P.lw = work; P.ls = work + n; P.lpr = work + 2 * n; P.un = work + 3 * n; P.ue = work + 4 * n; P.res = work + 5 * n;
  P.prog = prog; P.partial = partial; P.info = info; P.rsm_out = rsm_out;
  const int nbands = (jmax + 31) / 32;
  // the band-to-band rows live behind the progress counters (cfd_sip_sync_bytes); all entries start empty
  const int nb_alloc = (nrows - 1 + 31) / 32;
  P.brow = (double*)(prog + nb_alloc + 2);
  CFD_CUDA(cudaMemsetAsync(P.brow, 0xff, (size_t)nb_alloc * ld * sizeof(double), stream));
  const size_t smem = (size_t)SIP_W * SIP_WARP_DOUBLES * sizeof(double);
  CFD_CUDA(cudaFuncSetAttribute(k_sip_solve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0, nsm = 0;
  CFD_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_sip_solve, 32 * SIP_WARPS, smem));
  CFD_CUDA(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, device));
  // every SM takes part (the residual phase is a plain streaming pass); bands beyond nbands simply do not exist
  int grid = per_sm * nsm;
  (void)nbands;
  if (grid > max_partials) grid = max_partials;
  if (grid < 1) grid = 1;
  CFD_CUDA(cudaMemsetAsync(work, 0, 6 * n * sizeof(double), stream));
  CFD_CUDA(cudaMemsetAsync(prog, 0, (size_t)(nbands + 1) * sizeof(unsigned long long), stream));
  CFD_CUDA(cudaMemsetAsync(info, 0, 2 * sizeof(int), stream));
  void* args[] = {(void*)&P};
  CFD_CUDA(cudaLaunchCooperativeKernel((void*)k_sip_solve, dim3(grid), dim3(32 * SIP_WARPS), args, smem, stream));
  return CFD_OK;
}


// ---- sipsol as a drop-in on HOST arrays (include/cfd_b200.h: cfd_sipsol) ------------------------------------------------
// One call = one `call sipsol(i1, i2, j1, j2, aw, ae, as, an, ap, su, phi)` of fdm_implicit_cavity.f90:29-92 /
// fdm_explicit_cavity.f90:24-87 with i1 = j1 = 1: the six coefficient arrays and phi are uploaded, the whole solve (incomplete
// factorisation + up to maxit iterations) runs in one kernel, phi comes back.  Work space is allocated and freed per call:
// a convenience for hosts that keep their own time loop; the cfd_fdcav_* entry points keep everything resident.
extern "C" int cfd_sipsol(int device, int i2, int j2, int ld, const double* aw, const double* ae, const double* as_,
                          const double* an, const double* ap, const double* su, double* phi, int negate, double alpha,
                          double tol, int maxit, int* iters, double* rsm) {
  CFD_REQUIRE(i2 >= 3 && j2 >= 3 && ld >= i2 && aw && ae && as_ && an && ap && su && phi, "bad argument");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cfd_set_error("no CUDA device");
    return CFD_ERR_NO_DEVICE;
  }
  if (device < 0) CFD_CUDA(cudaGetDevice(&device));
  CFD_CUDA(cudaSetDevice(device));
  const size_t n = (size_t)ld * j2, bytes = n * sizeof(double);
  double* dev = nullptr;                       // 7 uploaded arrays + 6 work arrays + partials + rsm
  unsigned long long* sync = nullptr;
  int* info = nullptr;
  const size_t tail = 1024 + 8;
  if (cudaMalloc(&dev, 13 * bytes + tail * sizeof(double)) != cudaSuccess || cudaMalloc(&sync, cfd_sip_sync_bytes(ld, j2)) != cudaSuccess ||
      cudaMalloc(&info, 2 * sizeof(int)) != cudaSuccess) {
    if (dev) cudaFree(dev);
    if (sync) cudaFree(sync);
    if (info) cudaFree(info);
    cfd_set_error("cfd_sipsol: cudaMalloc of %zu bytes failed", 13 * bytes);
    return CFD_ERR_NOMEM;
  }
  const double* host[7] = {aw, ae, as_, an, ap, su, phi};
  int rc = CFD_OK;
  for (int k = 0; k < 7 && rc == CFD_OK; ++k)
    if (cudaMemcpy(dev + k * n, host[k], bytes, cudaMemcpyHostToDevice) != cudaSuccess) rc = CFD_ERR_CUDA;
  if (rc == CFD_OK) {
    const double* coef[5] = {dev, dev + n, dev + 2 * n, dev + 3 * n, dev + 4 * n};
    double* partial = dev + 13 * n;
    rc = sip_launch(device, 0, ld, j2, i2 - 2, j2 - 2, nullptr, coef, negate ? -1.0 : 1.0, alpha, tol, maxit, dev + 5 * n, dev + 6 * n,
                    dev + 7 * n, sync, partial, info, partial + 1024, 1024);
  }
  int hinfo[2] = {0, 0};
  if (rc == CFD_OK && (cudaMemcpy(phi, dev + 6 * n, bytes, cudaMemcpyDeviceToHost) != cudaSuccess ||
                       cudaMemcpy(hinfo, info, sizeof(hinfo), cudaMemcpyDeviceToHost) != cudaSuccess ||
                       (rsm && cudaMemcpy(rsm, dev + 13 * n + 1024, sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess)))
    rc = CFD_ERR_CUDA;
  cudaFree(dev);
  cudaFree(sync);
  cudaFree(info);
  if (rc != CFD_OK) {
    if (rc == CFD_ERR_CUDA) cfd_set_error("cfd_sipsol: %s", cudaGetErrorString(cudaGetLastError()));
    return rc;
  }
  if (hinfo[1]) {
    cfd_set_error("SIP: wavefront wait abandoned (pipeline stalled)");
    return CFD_ERR_CUDA;
  }
  if (iters) *iters = hinfo[0];
  return CFD_OK;
}
