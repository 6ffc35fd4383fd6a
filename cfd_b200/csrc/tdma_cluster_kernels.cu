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

#define XCH_N 9   // per block and line: E = (r, pL, pR) of chunk-end 0, A = (r, pL, pR) of chunk-end Tc-1, first node (cc0, cf0, ca0)

template <int M, int NC, int CS>
__global__ void __launch_bounds__(384, 1) k_tdma_flowc(const FlowParams P, const int S) {
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
  double* xch = bufB + 5 * Tc;                   // [2][XCH_N] interface numbers of this block, by line parity
  double* xlr = xch + 2 * XCH_N;                 // [2] XL, XR of the current line
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(xlr + 2);
  const unsigned full0 = smem_u32(bars), empty0 = smem_u32(bars + S), xbar = smem_u32(bars + 2 * S);
  int* counters = reinterpret_cast<int*>(bars + 2 * S + 1);
  const unsigned acq32 = smem_u32(counters), sto32 = smem_u32(counters + 1);
  const int ncw = Tc / 32;
  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, ncw);
    }
    mbar_init(xbar, CS);
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
    // ---------------- producer warp ---------------------------------------------------------------------
    if (tid == Tc) {
      const unsigned part_bytes = (unsigned)(Tc * sizeof(double));
      unsigned rs = 0, rph = 0;
      FlowIter it;
      int n = 0;
      for (it.init(P, cid, ncl); !it.done(P); it.next(P), ++n) {
        const bool zeroEW = it.p < P.n_zero_ew;
        const unsigned tx = part_bytes * ((unsigned)nc + (zeroEW ? 0u : 2u));
        const int i0 = it.line(P);
        while (lds_acquire(acq32) - (n + 1) < 0) __nanosleep(20);
        if (it.p > 0) asm volatile("fence.proxy.async;" ::: "memory");
        const double* csrc = P.coef + (long long)i0 * P.L.cSL + crank * Tc;
        const double* esrc = P.phi + (long long)(i0 + 1) * LD + crank * Tc;
        const double* wsrc = P.phi + (long long)(i0 - 1) * LD + crank * Tc;
        for (int k = 0; k < M; ++k) {
          const unsigned s = rs, ph = rph;
          if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
          mbar_wait(empty0 + 8 * s, ph ^ 1u);
          const unsigned fb = full0 + 8 * s;
          const unsigned dst = smem_u32(ring + (size_t)s * SW);
          mbar_expect_tx(fb, tx);
#pragma unroll
          for (int a = 0; a < nc; ++a) tma_load_1d(dst + a * part_bytes, csrc + (size_t)k * P.L.cSK + (size_t)a * T, part_bytes, fb);
          if (!zeroEW) {
            tma_load_1d(dst + nc * part_bytes, esrc + (size_t)k * T, part_bytes, fb);
            tma_load_1d(dst + (nc + 1) * part_bytes, wsrc + (size_t)k * T, part_bytes, fb);
          }
        }
      }
      if (blockIdx.x == 0) {
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
    unsigned rs = 0, rph = 0, xph = 0;
    bool peer_open[2] = {false, false};
    FlowIter it;
    it.init(P, cid, ncl);
    double bS = 0.0, bN = 0.0;
    if (!it.done(P) && !(it.p < P.n_zero_ends)) {
      const double* ph = P.phi + (long long)it.line(P) * LD;
      bS = __ldg(ph + q_s);
      bN = __ldg(ph + q_n);
    }
    for (int n = 0; !it.done(P); ++n) {
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
#pragma unroll
      for (int k = 0; k < M; ++k) {
        const unsigned s = rs, ph = rph;
        if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
        mbar_wait(full0 + 8 * s, ph);
        const unsigned sl = ring32 + s * stage_bytes;
        double a = lds_f64(sl), b = lds_f64(sl + T8), d = 1.0;
        unsigned sr = sl + 2 * T8;
        if (nc == 6) {
          d = lds_f64(sr);
          sr += T8;
        }
        double cv = lds_f64(sr + 2 * T8);
        if (!zeroEW) {
          cv = fma(lds_f64(sr + T8), lds_f64(sr + 4 * T8), cv);
          cv = fma(lds_f64(sr), lds_f64(sr + 3 * T8), cv);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty0 + 8 * s);
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
#pragma unroll
      for (int k = M - 3; k >= 0; --k) {
        cc[k] = fma(ca[k], cc[k + 1], cc[k]);
        cf[k] = fma(ca[k], cf[k + 1], cf[k]);
        ca[k] = ca[k] * ca[k + 1];
      }
      // ---- chunk-end relations  x_t = l*x_{t-1} + u*x_{t+1} + r + pL*XL + pR*XR
      double* xc = xch + (n & 1) * XCH_N;
      double l, u, r, pL = 0.0, pR = 0.0;
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
      }
      if (t == 0) {                       // first chunk of the block: its x_{t-1} is XL
        pL = l;
        l = 0.0;
        // first node of the block: x_first = cc0 + cf0*XL + ca0*x_end(0); with one node per chunk it IS x_end(0)
        xc[6] = (M == 1) ? 0.0 : cc[0];
        xc[7] = (M == 1) ? 0.0 : cf[0];
        xc[8] = (M == 1) ? 1.0 : ca[0];
      }
      double* cur = bufA;
      double* nxt = bufB;
      cur[t] = l;
      cur[Tc + t] = u;
      cur[2 * Tc + t] = r;
      cur[3 * Tc + t] = pL;
      cur[4 * Tc + t] = pR;
      consumer_sync(Tc);
      for (int sft = 1; sft < Tc; sft <<= 1) {
        double lm = 0.0, um = 0.0, rm = 0.0, am = 0.0, bm = 0.0, lp = 0.0, up = 0.0, rp = 0.0, ap = 0.0, bp = 0.0;
        if (t - sft >= 0) {
          lm = cur[t - sft];
          um = cur[Tc + t - sft];
          rm = cur[2 * Tc + t - sft];
          am = cur[3 * Tc + t - sft];
          bm = cur[4 * Tc + t - sft];
        }
        if (t + sft < Tc) {
          lp = cur[t + sft];
          up = cur[Tc + t + sft];
          rp = cur[2 * Tc + t + sft];
          ap = cur[3 * Tc + t + sft];
          bp = cur[4 * Tc + t + sft];
        }
        const double inv = rcp_nr(fma(-l, um, fma(-u, lp, 1.0)));
        r = fma(l, rm, fma(u, rp, r)) * inv;
        pL = fma(l, am, fma(u, ap, pL)) * inv;
        pR = fma(l, bm, fma(u, bp, pR)) * inv;
        l = l * lm * inv;
        u = u * up * inv;
        nxt[t] = l;
        nxt[Tc + t] = u;
        nxt[2 * Tc + t] = r;
        nxt[3 * Tc + t] = pL;
        nxt[4 * Tc + t] = pR;
        consumer_sync(Tc);
        double* tmp = cur;
        cur = nxt;
        nxt = tmp;
      }
      // ---- the one exchange of the line: interface relations of every block, then XL / XR of this block
      if (t == 0) {
        xc[0] = r;
        xc[1] = pL;
        xc[2] = pR;
      }
      if (t == Tc - 1) {
        xc[3] = r;
        xc[4] = pL;
        xc[5] = pR;
      }
      consumer_sync(Tc);
      if (t == 0) {
#pragma unroll
        for (int b = 0; b < CS; ++b) mbar_arrive_cluster(map_shared(xbar, (unsigned)b));
      }
      mbar_wait_cluster(xbar, xph);
      xph ^= 1u;
      if (t == 0) {
        // E_c = RE + LE*A_{c-1} + UE*F_{c+1};  A_c = RA + LA*A_{c-1} + UA*F_{c+1};  F_c = cc0 + cf0*A_{c-1} + ca0*E_c
        double RA[CS], LA[CS], UA[CS], PF[CS], QF[CS], SF[CS];
        const unsigned xc32 = smem_u32(xc);
#pragma unroll
        for (int b = 0; b < CS; ++b) {
          const unsigned ra = map_shared(xc32, (unsigned)b);
          const double RE = ld_dsmem(ra), LE = ld_dsmem(ra + 8), UE = ld_dsmem(ra + 16);
          RA[b] = ld_dsmem(ra + 24);
          LA[b] = ld_dsmem(ra + 32);
          UA[b] = ld_dsmem(ra + 40);
          const double c0 = ld_dsmem(ra + 48), f0 = ld_dsmem(ra + 56), a0 = ld_dsmem(ra + 64);
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
          // F_b = PF + QF*A_{b-1} + SF*F_{b+1},  A_{b-1} = al + be*F_b
          const double inv = rcp_nr(fma(-QF[b], be[b - 1], 1.0));
          ga[b] = fma(QF[b], al[b - 1], PF[b]) * inv;
          de[b] = SF[b] * inv;
          // A_b = RA + LA*A_{b-1} + UA*F_{b+1},  A_{b-1} = al + be*(ga + de*F_{b+1})
          al[b] = fma(LA[b], fma(be[b - 1], ga[b], al[b - 1]), RA[b]);
          be[b] = fma(LA[b] * be[b - 1], de[b], UA[b]);
        }
        double F[CS + 1], A[CS];
        F[CS] = 0.0;
#pragma unroll
        for (int b = CS - 1; b >= 1; --b) F[b] = fma(de[b], F[b + 1], ga[b]);
#pragma unroll
        for (int b = 0; b < CS; ++b) A[b] = fma(be[b], F[b + 1], al[b]);
        double XL = 0.0, XR = 0.0;
#pragma unroll
        for (int b = 0; b < CS; ++b) {
          if (b == crank) {
            XL = b > 0 ? A[b > 0 ? b - 1 : 0] : 0.0;
            XR = F[b + 1];
          }
        }
        xlr[0] = XL;
        xlr[1] = XR;
      }
      consumer_sync(Tc);
      const double XL = xlr[0], XR = xlr[1];
      const double xe = fma(pL, XL, fma(pR, XR, r));
      double xp;
      if (t > 0) xp = fma(cur[3 * Tc + t - 1], XL, fma(cur[4 * Tc + t - 1], XR, cur[2 * Tc + t - 1]));
      else xp = XL;
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
      consumer_sync(Tc);
      if (t == 0) sts_release(sto32, n + 1);
      bS = bS_next;
      bN = bN_next;
    }
  }
  __syncwarp();
  cluster_sync_all();   // no block leaves while another may still read its shared memory or signal its barrier
}

template <int M, int NC, int CS>
static cudaError_t launch_flowc_cs(cfd_ctx* c, const FlowParams& P, int S, size_t smem, int nmax) {
  const int Tc = P.L.T / CS;
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
  e = cudaLaunchKernelEx(&cfg, kern, P, S);
  if (e != cudaSuccess) {           // cooperative + cluster not accepted: the occupancy bound above still holds on an idle GPU
    (void)cudaGetLastError();
    cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, kern, P, S);
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
  const size_t fixed = (size_t)10 * Tc * sizeof(double) + (2 * XCH_N + 2) * sizeof(double) + 8 + 16;
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
