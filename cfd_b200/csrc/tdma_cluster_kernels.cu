// tdma_cluster_kernels.cu -- the flag-ordered line sweeps for lines too long for one thread block: a line is split over
// a thread-block CLUSTER of CS blocks (2, 4 or 8), each solving T/CS chunks of it.
//
// lisolv solves one tridiagonal system per S-N line (cavity_fvm.f90:62-80).  k_tdma_flow gives each line to one block of T
// threads, T*M >= nj; the register budget stops at T = 320 (M <= 16 nodes per thread), i.e. lines of 5120 nodes.  The
// 8192^2 configuration has 8194-node lines (T = 640) and north_star's 16384^2 target 16386-node lines (T = 1280).  Here:
//   * block c of the cluster owns chunks c*Tc .. (c+1)*Tc-1 (Tc = T/CS <= 320) and streams only its part of every slice
//     (nc+2 bulk copies of Tc doubles per slice, this GPU's memory only) through its own ring;
//   * inside a block the chunk-end unknowns are reduced by parallel cyclic reduction as before, but with two extra
//     right-hand sides: every unknown comes out as  x_t = r_t + pL_t*XL + pR_t*XR, XL = the last unknown of the block
//     before, XR = the first node of the block after (three-column PCR; 5 doubles per unknown);
//   * ONE exchange per line through distributed shared memory: each block publishes 9 numbers (the relations of its first
//     and last chunk-end unknown and of its first node), signals an mbarrier in every block of the cluster (remote
//     arrive), waits for its own, reads the others' numbers with ld.shared::cluster and solves the 2*CS-1 interface
//     unknowns redundantly by a short recurrence; then back-substitution and stores as in k_tdma_flow.
//   * a line is published (flag = epoch) by whichever block's helper lane arrives last at the line's arrival counter.
// Producer / helper roles, flags, peer halos and the work order are those of k_tdma_flow (flow.cuh).
#include <cuda.h>
#include "flow.cuh"

namespace cfd {

__device__ __forceinline__ unsigned cluster_rank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned cluster_id_x() {
  unsigned r;
  asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned ncluster_x() {
  unsigned r;
  asm volatile("mov.u32 %0, %%nclusterid.x;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned map_shared(unsigned laddr, unsigned rank) {
  unsigned r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(laddr), "r"(rank));
  return r;
}
__device__ __forceinline__ double ld_dsmem(unsigned raddr) {
  double v;
  asm volatile("ld.shared::cluster.f64 %0, [%1];" : "=d"(v) : "r"(raddr) : "memory");
  return v;
}
__device__ __forceinline__ void mbar_arrive_cluster(unsigned raddr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(raddr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(unsigned bar, unsigned parity) {
  unsigned ok;
  do {
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!ok);
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ int atom_add_acq_rel_gpu(int* p, int v) {
  int old;
  asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}

// asynchronous store of one double into another block's shared memory; completes `bytes` on that block's mbarrier
__device__ __forceinline__ void st_async_f64(unsigned raddr, double v, unsigned rbar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(raddr),
               "l"(__double_as_longlong(v)), "r"(rbar)
               : "memory");
}

// one tiled bulk copy (TMA) of a 3-D box out of this GPU's memory: [c2 rows][c1 segments][segment] -> dense shared memory
__device__ __forceinline__ void tma_load_3d(unsigned dst, const CUtensorMap* tm, int c0, int c1, int c2, unsigned bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
      "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
      : "memory");
}

#define XCH_N 9   // per block and line: E = (r, pL, pR) of chunk-end 0, A = (r, pL, pR) of chunk-end Tc-1, first node (cc0, cf0, ca0)

// TM: the nc coefficient parts of a slice arrive as ONE tiled copy through a tensor map of the coefficient block (rows of T
// doubles: row (i0*M + k)*nc + a; the block's Tc columns as Tc/SEG segments of SEG <= 256 doubles) instead of nc 1-D
// copies of Tc doubles.  Measured need: with seven copies of 2.5 KB per slice the copy engine, not HBM, set the pace
// (about 85 ns per copy).
template <int M, int NC, int CS>
__global__ void __launch_bounds__(384, 1) k_tdma_flowc(const FlowParams P, const int S, const int SEG,
                                                       const __grid_constant__ CUtensorMap tmap) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int T = P.L.T, Tc = T / CS;
  const long long LD = P.L.LD;
  const int tid = threadIdx.x;
  const int crank = (int)cluster_rank(), cid = (int)cluster_id_x(), ncl = (int)ncluster_x();
  constexpr int nc = NC;
  const int SW = (nc + 2) * Tc;
  double* ring = reinterpret_cast<double*>(smem_raw);
  double* bufA = ring + (size_t)S * SW;          // PCR: two buffers of 5*Tc
  double* bufB = bufA + 5 * Tc;
  double* xin = bufB + 5 * Tc;                   // [2][CS][XCH_N] interface numbers of every block of the cluster, by line parity
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(xin + 2 * CS * XCH_N);
  const unsigned full0 = smem_u32(bars), empty0 = smem_u32(bars + S), xbar0 = smem_u32(bars + 2 * S);   // xbar[2]
  int* counters = reinterpret_cast<int*>(bars + 2 * S + 2);
  const unsigned acq32 = smem_u32(counters), sto32 = smem_u32(counters + 1);
  const int ncw = Tc / 32;
  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, ncw);
    }
    mbar_init(xbar0, 1);
    mbar_init(xbar0 + 8, 1);
    counters[0] = 0;
    counters[1] = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (blockIdx.x == 0) {
      if (P.peer_ready[0]) st_release_sys(P.peer_ready[0], P.epoch0);
      if (P.peer_ready[1]) st_release_sys(P.peer_ready[1], P.epoch0);
    }
  }
  cluster_sync_all();   // every block's barriers exist before anyone signals them

  if (tid >= Tc + 32) {
    // ---------------- helper warp ---------------------------------------------------------------------------
    const int hl = tid - (Tc + 32);
    if (hl == 0) {
      FlowIter it;
      int n = 0;
      for (it.init(P, cid, ncl); !it.done(P); it.next(P), ++n) {
        if (it.p > 0) {
          const int i0 = it.line(P), oc = it.colour(P) ^ 1, want = P.epoch0 + it.p - 1;
          for (int nb = i0 - 1; nb <= i0 + 1; nb += 2) {
            if (nb < P.wait_lo[oc] || nb > P.wait_hi[oc]) continue;
            if (nb == P.halo_line[0] || nb == P.halo_line[1]) wait_flag<true>(P.flags + nb, want, P.abort_flag);
            else wait_flag<false>(P.flags + nb, want, P.abort_flag);
          }
        }
        sts_release(acq32, n + 1);
      }
    } else if (hl == 1) {
      FlowIter it;
      int n = 0;
      for (it.init(P, cid, ncl); !it.done(P); it.next(P), ++n) {
        while (lds_acquire(sto32) - (n + 1) < 0) __nanosleep(20);
        const int i0 = it.line(P), epoch = P.epoch0 + it.p;
        const int side = (i0 == P.edge_line[0]) ? 0 : ((i0 == P.edge_line[1]) ? 1 : -1);
        const bool peer = side >= 0 && P.peer_phi[side];
        if (peer) __threadfence_system();
        // the block that arrives last has (acquire) seen the others' parts and publishes the whole line
        const int old = atom_add_acq_rel_gpu(P.arrive + i0, 1);
        if ((old + 1) % CS == 0) {
          if (peer) {
            __threadfence_system();
            st_release_sys(P.peer_flag[side], epoch);
          }
          st_release_gpu(P.flags + i0, epoch);
        }
      }
    }
  } else if (tid >= Tc) {
    // ---------------- producer warp: lane a issues the copy of array a (nc coefficient parts, phi east, phi west) --------
    // (a block's part of a slice is nc+2 separate pieces of Tc doubles; issued one after the other by a single thread
    // the copies, not HBM, set the pace)
    {
      const int la = tid - Tc;
      const unsigned part_bytes = (unsigned)(Tc * sizeof(double));
      unsigned rs = 0, rph = 0;
      FlowIter it;
      int n = 0;
      for (it.init(P, cid, ncl); !it.done(P); it.next(P), ++n) {
        const bool zeroEW = it.p < P.n_zero_ew;
        const unsigned tx = part_bytes * ((unsigned)nc + (zeroEW ? 0u : 2u));
        const int i0 = it.line(P);
        if (la == 0) {
          while (lds_acquire(acq32) - (n + 1) < 0) __nanosleep(20);
          if (it.p > 0) asm volatile("fence.proxy.async;" ::: "memory");
        }
        __syncwarp();
        const bool tm = SEG > 0;
        // lane's source for node slice 0: with the tensor map lane 0 takes all nc coefficient parts in one tiled copy
        const double* src = la < nc ? P.coef + (long long)i0 * P.L.cSL + (size_t)la * T + crank * Tc
                                    : P.phi + (long long)(la == nc ? i0 + 1 : i0 - 1) * LD + crank * Tc;
        const size_t kstride = la < nc ? (size_t)P.L.cSK : (size_t)T;
        const bool mine = la < nc ? !tm : (la < nc + 2 && !zeroEW);
        const int row0 = i0 * M * nc, seg0 = tm ? crank * (Tc / SEG) : 0;
        for (int k = 0; k < M; ++k) {
          const unsigned s = rs, ph = rph;
          if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
          const unsigned fb = full0 + 8 * s;
          const unsigned dst = smem_u32(ring + (size_t)s * SW);
          if (la == 0) {
            mbar_wait(empty0 + 8 * s, ph ^ 1u);
            mbar_expect_tx(fb, tx);
            if (tm) tma_load_3d(dst, &tmap, 0, seg0, row0 + k * nc, fb);
          }
          __syncwarp();
          if (mine) tma_load_1d(dst + la * part_bytes, src + (size_t)k * kstride, part_bytes, fb);
        }
      }
      if (la == 0 && blockIdx.x == 0) {
        for (int sd = 0; sd < 2; ++sd)
          if (P.halo_line[sd] >= 0) wait_flag(P.flags + P.halo_line[sd], P.halo_last[sd], P.abort_flag);
      }
    }
  } else {
    // ---------------- consumers ---------------------------------------------------------------------------
    const int t = tid, lane = tid & 31;
    const int tg = crank * Tc + t;                 // chunk number in the whole line
    const int j0base = tg * M;
    const unsigned T8 = (unsigned)Tc * 8u, stage_bytes = (unsigned)SW * 8u, ring32 = smem_u32(ring) + (unsigned)t * 8u;
    const int q_s = lpos(P.jlo - 1, M, T), q_n = lpos(P.jhi + 1, M, T);
    unsigned rs = 0, rph = 0;
    bool peer_open[2] = {false, false};
    FlowIter it;
    it.init(P, cid, ncl);
    double bS = 0.0, bN = 0.0;
    if (!it.done(P) && !(it.p < P.n_zero_ends)) {
      const double* ph = P.phi + (long long)it.line(P) * LD;
      bS = __ldg(ph + q_s);
      bN = __ldg(ph + q_n);
    }
    const bool tm = P.dbg != nullptr && blockIdx.x == 0 && t == 0;
    unsigned long long tacc[6] = {0, 0, 0, 0, 0, 0}, nl_done = 0;
    for (int n = 0; !it.done(P); ++n) {
      long long c0 = tm ? clock64() : 0, c1;
#define STAMP(k)                \
  if (tm) {                     \
    c1 = clock64();             \
    tacc[k] += (unsigned long long)(c1 - c0); \
    c0 = c1;                    \
  }
      const bool zeroEW = it.p < P.n_zero_ew;
      const int i0 = it.line(P);
      double* phi = P.phi + (long long)i0 * LD;
      it.next(P);
      double bS_next = 0.0, bN_next = 0.0;
      if (!it.done(P) && !(it.p < P.n_zero_ends)) {
        const double* ph = P.phi + (long long)it.line(P) * LD;
        bS_next = __ldg(ph + q_s);
        bN_next = __ldg(ph + q_n);
      }
      double ca[M], cc[M], cf[M];
      // The elimination is a dependent chain per thread (reciprocal of the pivot, then the row): with lines this long
      // the consumers, not HBM, set the pace, so the pivot operands `as` (and `ap`) of slice k+1 are fetched -- and its
      // barrier waited for -- while slice k is still in the chain.
      unsigned s_cur = rs, ph_cur = rph;
      if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
      mbar_wait(full0 + 8 * s_cur, ph_cur);
      double b_nx = lds_f64(ring32 + s_cur * stage_bytes + T8), d_nx = (nc == 6) ? lds_f64(ring32 + s_cur * stage_bytes + 2 * T8) : 1.0;
#pragma unroll
      for (int k = 0; k < M; ++k) {
        const unsigned sl = ring32 + s_cur * stage_bytes;
        double a = lds_f64(sl), b = b_nx, d = d_nx;
        const unsigned sr = sl + (nc == 6 ? 3u : 2u) * T8;
        double cv = lds_f64(sr + 2 * T8);
        if (!zeroEW) {
          cv = fma(lds_f64(sr + T8), lds_f64(sr + 4 * T8), cv);
          cv = fma(lds_f64(sr), lds_f64(sr + 3 * T8), cv);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty0 + 8 * s_cur);
        if (k + 1 < M) {
          s_cur = rs;
          ph_cur = rph;
          if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
          mbar_wait(full0 + 8 * s_cur, ph_cur);
          b_nx = lds_f64(ring32 + s_cur * stage_bytes + T8);
          if (nc == 6) d_nx = lds_f64(ring32 + s_cur * stage_bytes + 2 * T8);
        }
        const int j0 = j0base + k;
        if (!(j0 >= P.jlo && j0 <= P.jhi)) {
          a = 0.0;
          b = 0.0;
          d = 1.0;
          cv = (j0 == P.jlo - 1) ? bS : ((j0 == P.jhi + 1) ? bN : 0.0);
        }
        if (k == 0) {
          const double inv = rcp_nr(d);
          ca[0] = a * inv;
          cc[0] = cv * inv;
          cf[0] = b * inv;
        } else {
          const double inv = rcp_nr(fma(-b, ca[k > 0 ? k - 1 : 0], d));
          ca[k] = a * inv;
          cc[k] = fma(b, cc[k > 0 ? k - 1 : 0], cv) * inv;
          cf[k] = b * cf[k > 0 ? k - 1 : 0] * inv;
        }
      }
      STAMP(0)
#pragma unroll
      for (int k = M - 3; k >= 0; --k) {
        cc[k] = fma(ca[k], cc[k + 1], cc[k]);
        cf[k] = fma(ca[k], cf[k + 1], cf[k]);
        ca[k] = ca[k] * ca[k + 1];
      }
      STAMP(1)
      // ---- chunk-end relations  x_t = l*x_{t-1} + u*x_{t+1} + r + pL*XL + pR*XR
      const int q = n & 1;
      double* xq = xin + (size_t)q * CS * XCH_N;                 // this line's table [CS][9]
      const unsigned xb = xbar0 + 8u * (unsigned)q;
      double l, u, r, pL = 0.0, pR = 0.0;
      double f0c = 0.0, f0f = 0.0, f0a = 1.0;                    // first node of the block (thread 0): x_first = f0c + f0f*XL + f0a*x_end(0)
      if (M == 1) {
        l = cf[0];
        u = ca[0];
        r = cc[0];
        if (t == Tc - 1) {                // its northern neighbour is the first node of the next block: XR itself
          pR = u;
          u = 0.0;
        }
      } else {
        bufB[t] = cc[0];
        bufB[Tc + t] = cf[0];
        bufB[2 * Tc + t] = ca[0];
        consumer_sync(Tc);
        if (t + 1 < Tc) {
          const double ncc = bufB[t + 1], ncf = bufB[Tc + t + 1], ng = bufB[2 * Tc + t + 1];
          const double inv = rcp_nr(fma(-ca[M - 1], ncf, 1.0));
          l = cf[M - 1] * inv;
          u = ca[M - 1] * ng * inv;
          r = fma(ca[M - 1], ncc, cc[M - 1]) * inv;
        } else {                          // last chunk of the block: x_last = cc + cf*x_prev + ca*XR
          l = cf[M - 1];
          u = 0.0;
          r = cc[M - 1];
          pR = ca[M - 1];
        }
        f0c = cc[0];
        f0f = cf[0];
        f0a = ca[0];
      }
      if (t == 0) {                       // first chunk of the block: its x_{t-1} is XL
        pL = l;
        l = 0.0;
      }
      // cyclic reduction with three right-hand sides; (l,u) and (r,pL) travel as 16-byte pairs
      double2* LU = reinterpret_cast<double2*>(bufA);            // two buffers of [Tc] double2 + [Tc] double2 + [Tc] double
      double2* RP = LU + Tc;
      double* QR = reinterpret_cast<double*>(RP + Tc);
      double2* LU2 = reinterpret_cast<double2*>(bufB);
      double2* RP2 = LU2 + Tc;
      double* QR2 = reinterpret_cast<double*>(RP2 + Tc);
      LU[t] = make_double2(l, u);
      RP[t] = make_double2(r, pL);
      QR[t] = pR;
      consumer_sync(Tc);
      for (int sft = 1; sft < Tc; sft <<= 1) {
        double2 lum = make_double2(0.0, 0.0), rpm = lum, lup = lum, rpp = lum;
        double qm = 0.0, qp = 0.0;
        if (t - sft >= 0) {
          lum = LU[t - sft];
          rpm = RP[t - sft];
          qm = QR[t - sft];
        }
        if (t + sft < Tc) {
          lup = LU[t + sft];
          rpp = RP[t + sft];
          qp = QR[t + sft];
        }
        const double inv = rcp_nr(fma(-l, lum.y, fma(-u, lup.x, 1.0)));
        r = fma(l, rpm.x, fma(u, rpp.x, r)) * inv;
        pL = fma(l, rpm.y, fma(u, rpp.y, pL)) * inv;
        pR = fma(l, qm, fma(u, qp, pR)) * inv;
        l = l * lum.x * inv;
        u = u * lup.y * inv;
        LU2[t] = make_double2(l, u);
        RP2[t] = make_double2(r, pL);
        QR2[t] = pR;
        consumer_sync(Tc);
        double2* t2 = LU; LU = LU2; LU2 = t2;
        t2 = RP; RP = RP2; RP2 = t2;
        double* t1 = QR; QR = QR2; QR2 = t1;
      }
      STAMP(2)
      // ---- the one exchange of the line.  Thread 0 holds E = (r,pL,pR) of chunk-end 0 and the first-node relation,
      // thread Tc-1 holds A = (r,pL,pR) of the last chunk-end: both write them into this block's row of the table and send
      // them to the same row in every other block with asynchronous shared-memory stores that complete on that block's
      // barrier -- no fence, no flag; each block expects (CS-1)*9 doubles per line.
      if (t == 0 || t == Tc - 1) {
        double* mine = xq + crank * XCH_N;
        const unsigned m32 = smem_u32(mine);
        if (t == 0) {
          mbar_expect_tx(xb, (unsigned)((CS - 1) * XCH_N * sizeof(double)));
          mine[0] = r; mine[1] = pL; mine[2] = pR; mine[6] = f0c; mine[7] = f0f; mine[8] = f0a;
#pragma unroll
          for (int b = 0; b < CS; ++b) {
            if (b == crank) continue;
            const unsigned ra = map_shared(m32, (unsigned)b), rb = map_shared(xb, (unsigned)b);
            st_async_f64(ra, r, rb); st_async_f64(ra + 8, pL, rb); st_async_f64(ra + 16, pR, rb);
            st_async_f64(ra + 48, f0c, rb); st_async_f64(ra + 56, f0f, rb); st_async_f64(ra + 64, f0a, rb);
          }
        }
        if (t == Tc - 1) {
          mine[3] = r; mine[4] = pL; mine[5] = pR;
#pragma unroll
          for (int b = 0; b < CS; ++b) {
            if (b == crank) continue;
            const unsigned ra = map_shared(m32, (unsigned)b), rb = map_shared(xb, (unsigned)b);
            st_async_f64(ra + 24, r, rb); st_async_f64(ra + 32, pL, rb); st_async_f64(ra + 40, pR, rb);
          }
        }
      }
      consumer_sync(Tc);                           // this block's own row is complete
      mbar_wait(xb, (unsigned)((n >> 1) & 1));     // ... and the other blocks' rows have arrived
      double XL = 0.0, XR = 0.0;
      {
        // every thread solves the interface unknowns itself (a short recurrence on the 9*CS numbers, broadcast reads):
        // E_c = RE + LE*A_{c-1} + UE*F_{c+1};  A_c = RA + LA*A_{c-1} + UA*F_{c+1};  F_c = cc0 + cf0*A_{c-1} + ca0*E_c
        double RA[CS], LA[CS], UA[CS], PF[CS], QF[CS], SF[CS];
#pragma unroll
        for (int b = 0; b < CS; ++b) {
          const double* row = xq + b * XCH_N;
          const double RE = row[0], LE = row[1], UE = row[2];
          RA[b] = row[3];
          LA[b] = row[4];
          UA[b] = row[5];
          const double c0 = row[6], f0 = row[7], a0 = row[8];
          PF[b] = fma(a0, RE, c0);
          QF[b] = fma(a0, LE, f0);
          SF[b] = a0 * UE;
        }
        // forward: A_c = al[c] + be[c]*F_{c+1},  F_c = ga[c] + de[c]*F_{c+1}   (A_{-1} = 0, F_CS = 0)
        double al[CS], be[CS], ga[CS], de[CS];
        al[0] = RA[0];
        be[0] = UA[0];
        ga[0] = 0.0;
        de[0] = 0.0;
#pragma unroll
        for (int b = 1; b < CS; ++b) {
          const double inv = rcp_nr(fma(-QF[b], be[b - 1], 1.0));
          ga[b] = fma(QF[b], al[b - 1], PF[b]) * inv;
          de[b] = SF[b] * inv;
          al[b] = fma(LA[b], fma(be[b - 1], ga[b], al[b - 1]), RA[b]);
          be[b] = fma(LA[b] * be[b - 1], de[b], UA[b]);
        }
        double F[CS + 1], A[CS];
        F[CS] = 0.0;
#pragma unroll
        for (int b = CS - 1; b >= 1; --b) F[b] = fma(de[b], F[b + 1], ga[b]);
#pragma unroll
        for (int b = 0; b < CS; ++b) A[b] = fma(be[b], F[b + 1], al[b]);
#pragma unroll
        for (int b = 0; b < CS; ++b) {
          if (b == crank) {
            XL = b > 0 ? A[b > 0 ? b - 1 : 0] : 0.0;
            XR = F[b + 1];
          }
        }
      }
      STAMP(3)
      const double xe = fma(pL, XL, fma(pR, XR, r));
      double xp;
      if (t > 0) {
        const double2 rpn = RP[t - 1];
        xp = fma(rpn.y, XL, fma(QR[t - 1], XR, rpn.x));
      } else {
        xp = XL;
      }
      const int side = (i0 == P.edge_line[0]) ? 0 : ((i0 == P.edge_line[1]) ? 1 : -1);
      double* rphi = nullptr;
      if (side >= 0 && P.peer_phi[side]) {
        if (!peer_open[side]) {
          if (t == 0) wait_flag(P.ready + side, P.epoch0, P.abort_flag);
          consumer_sync(Tc);
          peer_open[side] = true;
        }
        rphi = P.peer_phi[side];
      }
#pragma unroll
      for (int k = 0; k < M; ++k) {
        const int j0 = j0base + k;
        if (j0 >= P.jlo && j0 <= P.jhi) {
          const double xk = (k == M - 1) ? xe : fma(cf[k], xp, fma(ca[k], xe, cc[k]));
          phi[k * T + tg] = xk;
          if (rphi) rphi[k * T + tg] = xk;
        }
      }
      STAMP(4)
      consumer_sync(Tc);
      STAMP(5)
      ++nl_done;
      if (t == 0) sts_release(sto32, n + 1);
      bS = bS_next;
      bN = bN_next;
    }
#undef STAMP
    if (tm) {
      atomicAdd(P.dbg, nl_done);
      for (int k = 0; k < 6; ++k) atomicAdd(P.dbg + 1 + k, tacc[k]);
    }
  }
  __syncwarp();
  cluster_sync_all();   // no block leaves while another may still read its shared memory or signal its barrier
}

// Tensor map of a context's coefficient block, for the tiled copies of k_tdma_flowc: 3-D view (SEG, T/SEG, ni*M*nc rows) of
// doubles, box (SEG, Tc/SEG, nc).  Returns the segment length, 0 if the driver entry point is not available (the kernel
// then uses 1-D copies).
static int coef_tensor_map(cfd_ctx* c, int Tc, CUtensorMap* tm) {
  typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static encode_fn enc = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* fp = nullptr;
    cudaDriverEntryPointQueryResult qr;
    const char* e = getenv("CFD_B200_TDMA_TENSORMAP");
    if (!(e && atoi(e) == 0) && cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &qr) == cudaSuccess &&
        qr == cudaDriverEntryPointSuccess)
      enc = (encode_fn)fp;
    (void)cudaGetLastError();
  }
  if (!enc) return 0;
  const int T = c->L.T, nc = c->L.nc;
  const int seg = Tc <= 256 ? Tc : Tc / 2;
  if (Tc % seg != 0 || seg > 256 || (seg * 8) % 16 != 0) return 0;
  const cuuint64_t gdim[3] = {(cuuint64_t)seg, (cuuint64_t)(T / seg), (cuuint64_t)c->L.ni * c->L.M * nc};
  const cuuint64_t gstr[2] = {(cuuint64_t)seg * 8, (cuuint64_t)T * 8};
  const cuuint32_t box[3] = {(cuuint32_t)seg, (cuuint32_t)(Tc / seg), (cuuint32_t)nc};
  const cuuint32_t est[3] = {1, 1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, (void*)c->coef, gdim, gstr, box, est, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? seg : 0;
}

template <int M, int NC, int CS>
static cudaError_t launch_flowc_cs(cfd_ctx* c, const FlowParams& P, int S, size_t smem, int nmax) {
  const int Tc = P.L.T / CS;
  CUtensorMap tmap;
  memset(&tmap, 0, sizeof(tmap));
  const int SEG = coef_tensor_map(c, Tc, &tmap);
  auto kern = k_tdma_flowc<M, NC, CS>;
  cudaError_t e = cfd_raise_smem(c, kern);
  if (e != cudaSuccess) return e;
  if (CS > 8) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg{};
  cfg.blockDim = dim3(Tc + 64);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = c->stream;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CS;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeCooperative;
  at[1].val.cooperative = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  // how many clusters can be resident at once: blocks wait for lines other blocks publish, so the grid must not exceed it
  int ncl = 0;
  cfg.gridDim = dim3(c->num_sms / CS * CS);
  e = cudaOccupancyMaxActiveClusters(&ncl, kern, &cfg);
  if (e != cudaSuccess) return e;
  if (ncl < 1) return cudaErrorLaunchOutOfResources;
  if (ncl > nmax) ncl = nmax;
  cfg.gridDim = dim3(ncl * CS);
  cfg.numAttrs = 2;
  e = cudaLaunchKernelEx(&cfg, kern, P, S, SEG, tmap);
  if (e != cudaSuccess) {           // cooperative + cluster not accepted: the occupancy bound above still holds on an idle GPU
    (void)cudaGetLastError();
    cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, kern, P, S, SEG, tmap);
  }
  return e;
}

template <int M>
static cudaError_t launch_flowc_m(cfd_ctx* c, const FlowParams& P, int cs, int S, size_t smem, int nmax) {
  const bool five = P.L.nc == 5;
  switch (cs) {
    case 2: return five ? launch_flowc_cs<M, 5, 2>(c, P, S, smem, nmax) : launch_flowc_cs<M, 6, 2>(c, P, S, smem, nmax);
    case 4: return five ? launch_flowc_cs<M, 5, 4>(c, P, S, smem, nmax) : launch_flowc_cs<M, 6, 4>(c, P, S, smem, nmax);
    case 8: return five ? launch_flowc_cs<M, 5, 8>(c, P, S, smem, nmax) : launch_flowc_cs<M, 6, 8>(c, P, S, smem, nmax);
  }
  return cudaErrorInvalidValue;
}

}  // namespace cfd

using namespace cfd;

// blocks per cluster for a layout of T threads per line: 0 = this kernel does not apply
int cfd_cluster_size(int T) {
  if (T <= 320) return 0;
  for (int cs = 2; cs <= 8; cs *= 2)
    if (T % (32 * cs) == 0 && T / cs <= 320 && T / cs >= 64) return cs;
  return 0;
}

int cfd_launch_flow_cluster(cfd_ctx* c, FlowParams& P) {
  const int cs = cfd_cluster_size(c->L.T);
  if (!cs) {
    cfd_set_error("k_tdma_flowc: no cluster split for lines of %d layout threads", c->L.T);
    return CFD_ERR_UNSUPPORTED;
  }
  const int Tc = c->L.T / cs;
  P.arrive = c->flags + c->L.ni + CFD_FLAG_SLOTS;
  const size_t stage = (size_t)(c->L.nc + 2) * Tc * sizeof(double);
  const size_t fixed = (size_t)10 * Tc * sizeof(double) + (size_t)2 * cs * XCH_N * sizeof(double) + 16 + 16;
  int S = 16;
  const char* es = getenv("CFD_B200_TDMA_STAGES");
  if (es && atoi(es) > 0) S = atoi(es);
  while (S > 1 && S * stage + fixed + 2 * S * 8 > 220 * 1024) --S;
  const size_t smem = S * stage + fixed + 2 * (size_t)S * 8;
  const int nmax = P.nlines[0] > P.nlines[1] ? P.nlines[0] : P.nlines[1];
  cudaError_t e = cudaSuccess;
  switch (c->L.M) {
#define CFD_CASE(m) \
  case m:           \
    e = launch_flowc_m<m>(c, P, cs, S, smem, nmax); \
    break;
    CFD_CASE(4) CFD_CASE(5) CFD_CASE(6) CFD_CASE(7) CFD_CASE(8) CFD_CASE(9) CFD_CASE(10) CFD_CASE(11) CFD_CASE(12)
    CFD_CASE(13) CFD_CASE(14) CFD_CASE(15) CFD_CASE(16)
#undef CFD_CASE
    default:
      cfd_set_error("k_tdma_flowc: unsupported M=%d (long lines use 4..16 nodes per thread)", c->L.M);
      return CFD_ERR_INVALID;
  }
  if (e != cudaSuccess) {
    cfd_set_error("k_tdma_flowc: %s", cudaGetErrorString(e));
    return CFD_ERR_CUDA;
  }
  ++c->launches;
  return CFD_OK;
}
