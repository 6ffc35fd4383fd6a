// tdma_flow_kernels.cu -- all colour passes of an equation's line sweeps in ONE persistent launch, ordered by per-line
// flags instead of kernel boundaries, with the slab-edge lines handed to the neighbouring GPUs over peer memory.
//
// The reference runs `nswp` W->E sweeps of lisolv per equation (cavity_fvm.f90:189-191, 287-289, 343-345); with
// red/black lines that is 2*nswp colour passes.  k_tdma_tma (tdma_kernels.cu) runs one pass per launch: 18 launches per
// outer iteration, each starting with an empty ring and a cold HBM pipe, and -- on several GPUs -- each followed by a
// host-issued halo exchange.  Here:
//   * one launch walks through all passes; a line of pass p may start as soon as its two neighbour lines (other colour)
//     have finished pass p-1.  Every finished line publishes `flags[line] = epoch of the pass` (release); the producer
//     warp acquires the two neighbour flags before it streams the line's first slice.  No grid-wide barrier: blocks
//     drift from pass to pass and the copy ring never drains.
//   * i-slabs over several GPUs: the first / last owned line of a slab is also stored into the neighbour's halo line
//     (plain global stores into peer-mapped memory over NVLink), followed by a system-scope release of the neighbour's
//     flag for that halo line -- to the neighbour its halo line looks like any other line of the previous pass.  The edge
//     lines are solved FIRST in every pass, so the transfer has a whole pass to arrive.  Every bulk copy (TMA) of this
//     kernel reads this GPU's memory only; remote traffic is stores and flag polls.
//   * a neighbour may only store into a halo line once this GPU has launched its own sweep kernel (everything that read
//     the halo before is then complete in stream order): each launch first tells both neighbours "ready for epoch e0",
//     and the first remote store of a launch waits for the neighbour's ready mark.
// Same arithmetic as k_tdma_tma (same chunked elimination, same cyclic reduction): the two paths give identical bits.
// All waits are bounded; a timeout raises the context's abort flag instead of hanging the GPU.
#include "flow.cuh"

namespace cfd {

// Block = T consumer threads + a producer warp + a helper warp:
//   consumers        eliminate / reduce / back-substitute the line whose slices the ring delivers (as k_tdma_tma)
//   producer lane 0  issues the bulk copies (TMA) of the next slices as ring slots free up, across lines and passes
//   helper lane 0    "acquirer": for each coming line waits (acquire) for the flags of its two neighbour lines from the
//                    previous pass and tells the producer through a shared-memory counter -- the polls' latency stays
//                    off the producer's issue loop
//   helper lane 1    "publisher": once the consumers have stored a line (shared-memory counter, after a block barrier)
//                    it releases the line's flag at GPU scope and, for a slab-edge line, the neighbouring GPU's flag at
//                    system scope -- the fence that makes the stores visible costs about a microsecond, which nobody
//                    on the critical path waits for
template <int M, int BMAX, int NC>
__global__ void __launch_bounds__(BMAX, 1) k_tdma_flow(const FlowParams P, const int S) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int T = P.L.T;
  const long long LD = P.L.LD;
  const int tid = threadIdx.x;
  constexpr int nc = NC;
  const int SW = (nc + 2) * T;
  double* ring = reinterpret_cast<double*>(smem_raw);
  double* bufA = ring + (size_t)S * SW;
  double* bufB = bufA + 3 * T;
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(bufB + 3 * T);
  const unsigned full0 = smem_u32(bars), empty0 = smem_u32(bars + S);
  int* counters = reinterpret_cast<int*>(bars + 2 * S);      // [0] lines acquired, [1] lines stored
  const unsigned acq32 = smem_u32(counters), sto32 = smem_u32(counters + 1);
  const int ncw = T / 32;
  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, ncw);
    }
    counters[0] = 0;
    counters[1] = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    // tell both neighbours that every earlier reader of this GPU's halo lines is done (stream order): they may store
    if (blockIdx.x == 0) {
      if (P.peer_ready[0]) st_release_sys(P.peer_ready[0], P.epoch0);
      if (P.peer_ready[1]) st_release_sys(P.peer_ready[1], P.epoch0);
    }
  }
  __syncthreads();

  if (tid >= T + 32) {
    // ---------------- helper warp ---------------------------------------------------------------------------
    const int hl = tid - (T + 32);
    if (hl == 0) {
      // acquirer
      FlowIter it;
      int n = 0;
      for (it.init(P); !it.done(P); it.next(P), ++n) {
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
      // publisher
      FlowIter it;
      int n = 0;
      for (it.init(P); !it.done(P); it.next(P), ++n) {
        while (lds_acquire(sto32) - (n + 1) < 0) __nanosleep(20);
        const int i0 = it.line(P), epoch = P.epoch0 + it.p;
        const int side = (i0 == P.edge_line[0]) ? 0 : ((i0 == P.edge_line[1]) ? 1 : -1);
        if (side >= 0 && P.peer_phi[side]) {
          __threadfence_system();
          st_release_sys(P.peer_flag[side], epoch);
        }
        st_release_gpu(P.flags + i0, epoch);
      }
    }
    return;
  }

  if (tid >= T) {
    // ---------------- producer warp ---------------------------------------------------------------------
    if (tid == T) {
      const unsigned slice_bytes = (unsigned)(T * sizeof(double));
      unsigned rs = 0, rph = 0;
      FlowIter it;
      int n = 0;
      for (it.init(P); !it.done(P); it.next(P), ++n) {
        const bool zeroEW = it.p < P.n_zero_ew;
        const unsigned tx = slice_bytes * ((unsigned)nc + (zeroEW ? 0u : 2u));
        const int i0 = it.line(P);
        // the neighbour lines were rewritten by the previous pass: the acquirer has seen both complete
        while (lds_acquire(acq32) - (n + 1) < 0) __nanosleep(20);
        if (it.p > 0) asm volatile("fence.proxy.async;" ::: "memory");   // generic-proxy writes (other SMs / peer GPU) -> bulk-copy reads
        const double* csrc = P.coef + (long long)i0 * P.L.cSL;
        const double* esrc = P.phi + (long long)(i0 + 1) * LD;
        const double* wsrc = P.phi + (long long)(i0 - 1) * LD;
        for (int k = 0; k < M; ++k) {
          const unsigned s = rs, ph = rph;
          if (++rs == (unsigned)S) { rs = 0; rph ^= 1u; }
          mbar_wait(empty0 + 8 * s, ph ^ 1u);
          const unsigned fb = full0 + 8 * s;
          const unsigned dst = smem_u32(ring + (size_t)s * SW);
          mbar_expect_tx(fb, tx);
          tma_load_1d(dst, csrc + (size_t)k * P.L.cSK, nc * slice_bytes, fb);
          if (!zeroEW) {
            tma_load_1d(dst + nc * slice_bytes, esrc + (size_t)k * T, slice_bytes, fb);
            tma_load_1d(dst + (nc + 1) * slice_bytes, wsrc + (size_t)k * T, slice_bytes, fb);
          }
        }
      }
      // The neighbours' last stores into my halo lines must have landed before this launch ends (what follows on the
      // stream reads them).  Their edge lines are solved first in every pass, so this is normally already true.
      if (blockIdx.x == 0) {
        for (int sd = 0; sd < 2; ++sd)
          if (P.halo_line[sd] >= 0) wait_flag(P.flags + P.halo_line[sd], P.halo_last[sd], P.abort_flag);
      }
    }
    return;
  }

  // ---------------- consumers ---------------------------------------------------------------------------
  const int t = tid, lane = tid & 31;
  const int j0base = t * M;
  const unsigned T8 = (unsigned)T * 8u, stage_bytes = (unsigned)SW * 8u, ring32 = smem_u32(ring) + (unsigned)t * 8u;
  const int q_s = lpos(P.jlo - 1, M, T), q_n = lpos(P.jhi + 1, M, T);
  unsigned rs = 0, rph = 0;
  bool peer_open[2] = {false, false};   // this launch has seen the neighbour's ready mark
  FlowIter it;
  it.init(P);
  // Dirichlet end values phi(i, jstart-1), phi(i, nj) of a line are fetched one item ahead (never written by a sweep:
  // safe to read at any time)
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
    double l, u, r;
    if (M == 1) {
      l = cf[0];
      u = ca[0];
      r = cc[0];
    } else {
      bufB[t] = cc[0];
      bufB[T + t] = cf[0];
      bufB[2 * T + t] = ca[0];
      consumer_sync(T);
      double ncc = 0.0, ncf = 0.0, ng = 0.0;
      if (t + 1 < T) {
        ncc = bufB[t + 1];
        ncf = bufB[T + t + 1];
        ng = bufB[2 * T + t + 1];
      }
      const double inv = rcp_nr(fma(-ca[M - 1], ncf, 1.0));
      l = cf[M - 1] * inv;
      u = ca[M - 1] * ng * inv;
      r = fma(ca[M - 1], ncc, cc[M - 1]) * inv;
    }
    double* cur = bufA;
    double* nxt = bufB;
    cur[t] = l;
    cur[T + t] = u;
    cur[2 * T + t] = r;
    consumer_sync(T);
    for (int sft = 1; sft < T; sft <<= 1) {
      double lm = 0.0, um = 0.0, rm = 0.0, lp = 0.0, up = 0.0, rp = 0.0;
      if (t - sft >= 0) {
        lm = cur[t - sft];
        um = cur[T + t - sft];
        rm = cur[2 * T + t - sft];
      }
      if (t + sft < T) {
        lp = cur[t + sft];
        up = cur[T + t + sft];
        rp = cur[2 * T + t + sft];
      }
      const double inv = rcp_nr(fma(-l, um, fma(-u, lp, 1.0)));
      r = fma(l, rm, fma(u, rp, r)) * inv;
      l = l * lm * inv;
      u = u * up * inv;
      nxt[t] = l;
      nxt[T + t] = u;
      nxt[2 * T + t] = r;
      consumer_sync(T);
      double* tmp = cur;
      cur = nxt;
      nxt = tmp;
    }
    const double xe = r;
    const double xp = (t > 0) ? cur[2 * T + t - 1] : 0.0;
    // a slab-edge line also goes into the neighbour's halo line (peer memory)
    const int side = (i0 == P.edge_line[0]) ? 0 : ((i0 == P.edge_line[1]) ? 1 : -1);
    double* rphi = nullptr;
    if (side >= 0 && P.peer_phi[side]) {
      if (!peer_open[side]) {
        // the neighbour must have launched this same sweep kernel (its earlier readers of the halo are then done)
        if (t == 0) wait_flag(P.ready + side, P.epoch0, P.abort_flag);
        consumer_sync(T);
        peer_open[side] = true;
      }
      rphi = P.peer_phi[side];
    }
#pragma unroll
    for (int k = 0; k < M; ++k) {
      const int j0 = j0base + k;
      if (j0 >= P.jlo && j0 <= P.jhi) {
        const double xk = (k == M - 1) ? xe : fma(cf[k], xp, fma(ca[k], xe, cc[k]));
        phi[k * T + t] = xk;
        if (rphi) rphi[k * T + t] = xk;
      }
    }
    consumer_sync(T);   // every thread's stores precede the publisher's release; the PCR buffers are free again
    if (t == 0) sts_release(sto32, n + 1);
    bS = bS_next;
    bN = bN_next;
  }
}

// ---- halo push over peer memory: the lines a neighbour needs go straight into its halo lines --------------------------
// Handshake per call (xe = the context's exchange counter, the same on both sides because both run the same call
// sequence): (1) tell each neighbour "my earlier readers of the halo are done, you may write" (ready = xe), (2) wait
// for the neighbour's ready, (3) store the lines, system fence, (4) release "data = xe" at the neighbour, (5) wait for
// both neighbours' data marks.  One block.
__global__ void __launch_bounds__(1024, 1) k_halo_push(const HaloPush H) {
  __shared__ int ok;
  if (threadIdx.x == 0) {
    ok = 1;
    for (int s = 0; s < 2; ++s)
      if (H.peer_xready[s]) st_release_sys(H.peer_xready[s], H.xe);
    for (int s = 0; s < 2; ++s)
      if (H.peer_xready[s] && !wait_flag(H.xready + s, H.xe, H.abort_flag)) ok = 0;
  }
  __syncthreads();
  if (ok) {
    for (int m = 0; m < H.nmsg; ++m) {
      const double* __restrict__ src = H.src[m];
      double* __restrict__ dst = H.dst[m];
      const long long n2 = H.count[m] >> 1;    // LD is even (T multiple of 32): whole double2 transfers
      const double2* s2 = reinterpret_cast<const double2*>(src);
      double2* d2 = reinterpret_cast<double2*>(dst);
      for (long long q = threadIdx.x; q < n2; q += blockDim.x) d2[q] = s2[q];
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s)
      if (H.peer_xdata[s]) st_release_sys(H.peer_xdata[s], H.xe);
    for (int s = 0; s < 2; ++s)
      if (H.peer_xdata[s]) wait_flag(H.xdata + s, H.xe, H.abort_flag);
  }
}

// a slab that has no line to solve in this call still owes its neighbours the "ready" mark of the launch
__global__ void k_flow_ready(int* r0, int* r1, int epoch0) {
  if (r0) st_release_sys(r0, epoch0);
  if (r1) st_release_sys(r1, epoch0);
}

// ---- one-to-all broadcast of a scalar over peer memory --------------------------------------------------------------
// The rank that owns pp(ipref,jpref) stores the value into slot (epoch & 3) of every rank's mailbox and releases the
// rank's broadcast flag; every rank (the owner too) waits for its own flag before the correction kernel reads the slot.
// Four slots: a rank can be at most a few exchange stages behind the owner, never four broadcasts.
struct BcastParams {
  int n;
  const double* src;
  int* flags[CFD_BCAST_MAX];
  double* mail[CFD_BCAST_MAX];
  int epoch;
};
__global__ void k_scalar_send(const BcastParams B) {
  const int r = threadIdx.x;
  if (r >= B.n) return;
  B.mail[r][B.epoch & 3] = *B.src;
  __threadfence_system();
  st_release_sys(B.flags[r], B.epoch);
}
__global__ void k_scalar_wait(const int* flag, int epoch, int* abort_flag) { wait_flag<true>(flag, epoch, abort_flag); }

template <int M>
static cudaError_t launch_flow_m(cfd_ctx* c, const FlowParams& P, int nblocks, int S, size_t smem) {
  const int nthreads = P.L.T + 64;   // consumers + producer warp + helper warp
  void* args[2] = {(void*)&P, (void*)&S};
  cudaError_t e;
  // cooperative launch: every block must be resident, since blocks wait for lines other blocks publish
#define LAUNCH_FLOW(NCV)                                                                                        \
  {                                                                                                             \
    e = cfd_raise_smem(c, k_tdma_flow<M, 384, NCV>);                                                            \
    if (e == cudaSuccess)                                                                                       \
      e = cudaLaunchCooperativeKernel((const void*)k_tdma_flow<M, 384, NCV>, dim3(nblocks), dim3(nthreads), args, \
                                      smem, c->stream);                                                         \
  }
  if (P.L.nc == 5) LAUNCH_FLOW(5) else LAUNCH_FLOW(6)
#undef LAUNCH_FLOW
  return e;
}

}  // namespace cfd

using namespace cfd;

// True when cfd_launch_sweeps_flow can take this context: red/black order, lines in one piece (no j-slab), a layout
// that fits the 352-thread build of the TMA-staged solver.
bool cfd_flow_eligible(const cfd_ctx* c) {
  static int use = -1;
  if (use < 0) {
    const char* e = getenv("CFD_B200_FLOW");
    use = e ? atoi(e) : 1;
  }
  const char* tma = getenv("CFD_B200_TDMA_TMA");
  if (tma && atoi(tma) == 0) return false;
  // an i-slab needs its neighbours attached: without them the halo lines change between passes by host-issued
  // exchanges, which only the one-pass-per-launch path can interleave
  const bool peers_ok = (c->cfg.halo_w == 0 || c->peer_attached[0]) && (c->cfg.halo_e == 0 || c->peer_attached[1]);
  // Lines of 321 .. 992 layout threads have two solvers: one block per line (k_tdma_tma's 672 / 1024-thread builds, one
  // launch per colour pass) and a cluster of blocks per line (k_tdma_flowc).  Measured at 8192^2 (T = 640): three sweeps
  // 2.18 ms against 2.55 ms on one GPU, and the buoyancy iteration over two GPUs 8.07 ms (one block per line, NCCL halo
  // lines) against 8.32 ms (clusters, halo lines over peer memory): the cluster kernel's three-column reduction costs
  // more than it saves there, so those layouts keep the first solver.  Beyond 992 threads the cluster kernel is the
  // only one.  Narrow i-slabs turn the balance: with 2048 lines per GPU (8192^2 over 4) the same iteration takes 3.99 ms
  // with clusters and peer-memory halos against 4.48 ms with NCCL stages, with 1024 lines (over 8) 2.17 against 2.81 ms --
  // every colour pass of the one-block solver ends in a partial wave and in a host-issued exchange.  Slabs of at most 3072
  // local lines therefore take the cluster solver.  CFD_B200_CLUSTER=1 forces it everywhere (tests), =0 never prefers it.
  const char* fce = getenv("CFD_B200_CLUSTER");          // read per call: tests switch it
  const bool force_cluster = fce && atoi(fce) != 0;
  const bool narrow_slab = (c->cfg.halo_w > 0 || c->cfg.halo_e > 0) && c->L.ni <= 3072 && !(fce && atoi(fce) == 0);
  const bool cluster_ok = cfd_cluster_size(c->L.T) > 0 && c->L.M >= 4 && (c->L.T + 32 > 1024 || force_cluster || narrow_slab);
  const bool layout_ok = (c->L.T >= 128 && c->L.T <= 320) || cluster_ok;
  return use && c->cfg.order == CFD_ORDER_REDBLACK && c->cfg.halo_s == 0 && c->cfg.halo_n == 0 && c->cfg.joff == 0 &&
         layout_ok && c->flags != nullptr && peers_ok;
}

static int flow_launch(cfd_ctx* c, int istart, int jstart, int field, int npass, int colour0, int n_zero_ew, int n_zero_ends);

// nsweeps red/black sweeps (2*nsweeps colour passes) of `phi` over the current coefficients, one launch.
int cfd_launch_sweeps_flow(cfd_ctx* c, int istart, int jstart, int field, int nsweeps, int first) {
  if (nsweeps <= 0) return CFD_OK;
  return flow_launch(c, istart, jstart, field, 2 * nsweeps, 0, first ? 1 : 0, first ? 2 : 0);
}

// one colour pass through the same kernels (lines longer than one block have no other solver)
int cfd_launch_sweep_flow_pass(cfd_ctx* c, int istart, int jstart, int field, int colour, int first) {
  return flow_launch(c, istart, jstart, field, 1, colour, (first && colour == 0) ? 1 : 0, first ? 1 : 0);
}

static int flow_launch(cfd_ctx* c, int istart, int jstart, int field, int npass, int colour0, int n_zero_ew, int n_zero_ends) {
  const int epoch0 = c->epoch;
  c->epoch += npass;                // advanced on every rank alike, whether or not this slab has lines to solve
  const int NI = c->cfg.ni_global ? c->cfg.ni_global : c->L.ni;
  const int own_lo = c->cfg.goff + c->cfg.halo_w + 1, own_hi = c->cfg.goff + c->L.ni - c->cfg.halo_e;   // global, 1-based
  int glo = own_lo, ghi = own_hi;
  if (glo < istart) glo = istart;
  if (ghi > NI - 1) ghi = NI - 1;
  FlowParams P{};
  P.L = c->L;
  P.npass = npass;
  P.colour0 = colour0;
  for (int col = 0; col < 2; ++col) {
    const int g0 = glo + ((((glo - istart) & 1) != col) ? 1 : 0);
    if (g0 > ghi) {
      P.i_first[col] = 0;
      P.nlines[col] = 0;
    } else {
      P.i_first[col] = g0 - 1 - c->cfg.goff;
      P.nlines[col] = (ghi - g0) / 2 + 1;
    }
    // lines of this colour whose flag another pass may wait for: the owned ones, plus a halo line next to an owned
    // line when the neighbouring slab updates it (it lies inside the global sweep range istart .. NI-1)
    int lo = P.i_first[col], hi = P.i_first[col] + 2 * (P.nlines[col] - 1);
    if (P.nlines[col] == 0) { lo = 1; hi = 0; }
    const int gw = own_lo - 1, ge = own_hi + 1;                         // global numbers of the adjacent halo lines
    if (c->cfg.halo_w > 0 && c->peer_attached[0] && gw >= istart && ((gw - istart) & 1) == col) {
      const int l = gw - 1 - c->cfg.goff;
      if (P.nlines[col] == 0) { lo = hi = l; } else if (l < lo) lo = l;
    }
    if (c->cfg.halo_e > 0 && c->peer_attached[1] && ge <= NI - 1 && ((ge - istart) & 1) == col) {
      const int l = ge - 1 - c->cfg.goff;
      if (hi < lo) { lo = hi = l; } else if (l > hi) hi = l;
    }
    P.wait_lo[col] = lo;
    P.wait_hi[col] = hi;
  }
  int jlo, jhi;
  const int NJ = c->L.nj;
  cfd_local_j_range(c, jstart - 1, NJ - 2, &jlo, &jhi);
  if (P.nlines[0] + P.nlines[1] == 0 || jhi < jlo) {
    int* r[2] = {nullptr, nullptr};
    for (int s = 0; s < 2; ++s)
      if (c->peer_attached[s]) r[s] = c->peer_flags[s] + c->peer_ni[s] + CFD_FLAG_SWEEP_READY + (1 - s);
    if (r[0] || r[1]) {
      k_flow_ready<<<1, 1, 0, c->stream>>>(r[0], r[1], epoch0);
      ++c->launches;
      CFD_CUDA(cudaGetLastError());
    }
    return CFD_OK;
  }
  P.jlo = jlo;
  P.jhi = jhi;
  P.n_zero_ew = n_zero_ew;
  P.n_zero_ends = n_zero_ends;
  P.coef = c->field[CFD_AN];
  P.phi = c->field[field];
  P.flags = c->flags;
  P.epoch0 = epoch0;
  P.abort_flag = c->flags + c->L.ni + CFD_FLAG_ABORT;
  P.ready = c->flags + c->L.ni + CFD_FLAG_SWEEP_READY;
  for (int s = 0; s < 2; ++s) {
    P.halo_line[s] = -1;
    const int gh = s == 0 ? own_lo - 1 : own_hi + 1;       // the halo line next to my owned lines, global number
    if (c->peer_attached[s] && (s == 0 ? c->cfg.halo_w : c->cfg.halo_e) > 0 && gh >= istart && gh <= NI - 1) {
      const int ch = (gh - istart) & 1;                     // its colour; passes of that colour: p with (p & 1) == ch
      const int plast = ((P.npass - 1) & 1) == ch ? P.npass - 1 : P.npass - 2;
      P.halo_line[s] = gh - 1 - c->cfg.goff;
      P.halo_last[s] = epoch0 + plast;
    }
  }
  for (int s = 0; s < 2; ++s) {
    P.edge_line[s] = -1;
    P.peer_phi[s] = nullptr;
    P.peer_flag[s] = nullptr;
    P.peer_ready[s] = nullptr;
    if (!c->peer_attached[s] || !c->peer_field[s][field]) continue;
    // my first (s = 0) / last (s = 1) owned line, if the sweep updates it
    const int g = s == 0 ? own_lo : own_hi;
    const int peer_line = s == 0 ? c->peer_halo_line[0] : c->peer_halo_line[1];
    P.peer_ready[s] = c->peer_flags[s] + c->peer_ni[s] + CFD_FLAG_SWEEP_READY + (1 - s);   // I am its east (s=0) / west (s=1) neighbour
    if (g < glo || g > ghi) continue;
    P.edge_line[s] = g - 1 - c->cfg.goff;
    P.peer_phi[s] = c->peer_field[s][field] + (long long)peer_line * c->L.LD;
    P.peer_flag[s] = c->peer_flags[s] + peer_line;
  }
  {
    static int timing = -1;
    if (timing < 0) {
      const char* e = getenv("CFD_B200_FLOW_TIMING");
      timing = e ? atoi(e) : 0;
    }
    if (timing && !c->dbg) {
      CFD_CUDA(cudaMalloc(&c->dbg, 8 * sizeof(unsigned long long)));
      CFD_CUDA(cudaMemset(c->dbg, 0, 8 * sizeof(unsigned long long)));
    }
    P.dbg = c->dbg;
  }
  if (cfd_cluster_size(c->L.T) > 0) return cfd_launch_flow_cluster(c, P);    // a line split over a cluster of blocks
  const size_t stage = (size_t)(c->L.nc + 2) * c->L.T * sizeof(double), pcr = (size_t)6 * c->L.T * sizeof(double);
  int S = 16;
  const char* es = getenv("CFD_B200_TDMA_STAGES");
  if (es && atoi(es) > 0) S = atoi(es);
  while (S > 1 && S * stage + pcr + 2 * S * 8 + 16 > 220 * 1024) --S;
  const size_t smem = S * stage + pcr + 2 * (size_t)S * 8 + 16;
  int nmax = P.nlines[0] > P.nlines[1] ? P.nlines[0] : P.nlines[1];
  const int nblocks = nmax < c->num_sms ? nmax : c->num_sms;
  cudaError_t e = cudaSuccess;
  switch (c->L.M) {
#define CFD_CASE(m) \
  case m:           \
    e = launch_flow_m<m>(c, P, nblocks, S, smem); \
    break;
    CFD_CASE(1) CFD_CASE(2) CFD_CASE(3) CFD_CASE(4) CFD_CASE(5) CFD_CASE(6) CFD_CASE(7) CFD_CASE(8)
    CFD_CASE(9) CFD_CASE(10) CFD_CASE(11) CFD_CASE(12) CFD_CASE(13) CFD_CASE(14) CFD_CASE(15) CFD_CASE(16)
#undef CFD_CASE
    default:
      cfd_set_error("unsupported M=%d", c->L.M);
      return CFD_ERR_INVALID;
  }
  if (e != cudaSuccess) {
    cfd_set_error("k_tdma_flow: %s", cudaGetErrorString(e));
    return CFD_ERR_CUDA;
  }
  ++c->launches;
  return CFD_OK;
}

// Push halo lines of up to 4 fields to the neighbouring slabs: `w` lines to the east neighbour's west halo (my last
// owned lines), `e` lines to the west neighbour's east halo (my first owned lines); both sides wait for what they receive.
int cfd_launch_halo_push(cfd_ctx* c, int nfields, const int* fields, const int* w, const int* e) {
  HaloPush H{};
  const int ni = c->L.ni, hw = c->cfg.halo_w, he = c->cfg.halo_e, nown = ni - hw - he;
  const long long LD = c->L.LD;
  int* fl = c->flags + ni;
  H.xe = ++c->xepoch;
  H.abort_flag = fl + CFD_FLAG_ABORT;
  H.xready = fl + CFD_FLAG_XREADY;
  H.xdata = fl + CFD_FLAG_XDATA;
  for (int s = 0; s < 2; ++s) {
    if (!c->peer_attached[s]) continue;
    int* pf = c->peer_flags[s] + c->peer_ni[s];
    H.peer_xready[s] = pf + CFD_FLAG_XREADY + (1 - s);
    H.peer_xdata[s] = pf + CFD_FLAG_XDATA + (1 - s);
  }
  for (int k = 0; k < nfields; ++k) {
    const int f = fields[k];
    if (f < 0 || f >= CFD_FIELD_COUNT || !c->field[f] || cfd_is_coef_field(f)) {
      cfd_set_error("cfd_halo_push: field %d is not a plain field of this context", f);
      return CFD_ERR_INVALID;
    }
    if (c->peer_attached[0] && e[k] > 0) {            // my first owned lines -> west neighbour's east halo
      if (!c->peer_field[0][f]) { cfd_set_error("cfd_halo_push: field %d has no west peer mapping", f); return CFD_ERR_INVALID; }
      const int n = e[k];
      H.src[H.nmsg] = c->field[f] + (long long)hw * LD;
      H.dst[H.nmsg] = c->peer_field[0][f] + (long long)c->peer_halo_line[0] * LD;
      H.count[H.nmsg++] = n * LD;
    }
    if (c->peer_attached[1] && w[k] > 0) {            // my last owned lines -> east neighbour's west halo
      if (!c->peer_field[1][f]) { cfd_set_error("cfd_halo_push: field %d has no east peer mapping", f); return CFD_ERR_INVALID; }
      const int n = w[k] < c->peer_hw[1] ? w[k] : c->peer_hw[1];
      if (n > 0) {
        H.src[H.nmsg] = c->field[f] + (long long)(hw + nown - n) * LD;
        H.dst[H.nmsg] = c->peer_field[1][f] + (long long)(c->peer_hw[1] - n) * LD;
        H.count[H.nmsg++] = n * LD;
      }
    }
  }
  k_halo_push<<<1, 1024, 0, c->stream>>>(H);
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  return CFD_OK;
}

// One colour pass described by SweepParams (cfd_sweep / cfd_sweep_part) through the flag-ordered kernels: what
// launch_sweep (tdma_kernels.cu) falls back to when the lines are longer than any single-block solver takes.
int cfd_launch_flow_from_sweep(cfd_ctx* c, const cfd::SweepParams& Q, int field_unused) {
  (void)field_unused;
  if (Q.basis || Q.end_s_const || Q.end_n_const || Q.i_step != 2 || Q.out != Q.phi) {
    cfd_set_error("lines of %d layout threads: only plain colour passes are available (no j-slab pieces)", c->L.T);
    return CFD_ERR_UNSUPPORTED;
  }
  FlowParams P{};
  P.L = c->L;
  P.npass = 1;
  P.colour0 = 0;
  P.i_first[0] = Q.i_first;
  P.nlines[0] = Q.nlines;
  P.i_first[1] = 0;
  P.nlines[1] = 0;
  P.wait_lo[0] = P.wait_lo[1] = 1;
  P.wait_hi[0] = P.wait_hi[1] = 0;
  P.jlo = Q.jlo;
  P.jhi = Q.jhi;
  P.n_zero_ew = (Q.first && Q.first_colour == 0) ? 1 : 0;
  P.n_zero_ends = Q.first ? 1 : 0;
  P.coef = Q.an;
  P.phi = Q.phi;
  P.flags = c->flags;
  P.epoch0 = c->epoch;
  c->epoch += 1;
  P.abort_flag = c->flags + c->L.ni + CFD_FLAG_ABORT;
  P.ready = c->flags + c->L.ni + CFD_FLAG_SWEEP_READY;
  for (int s = 0; s < 2; ++s) {
    P.edge_line[s] = -1;
    P.halo_line[s] = -1;
  }
  if (Q.nlines <= 0) return CFD_OK;
  return cfd_launch_flow_cluster(c, P);
}

// owner != 0: this context holds element (i,j) (local, 1-based) of `field` and every rank's flag block is attached
// (cfd_bcast_attach).  *value_dev = device address of the received value, valid until three more broadcasts.
int cfd_launch_scalar_bcast(cfd_ctx* c, int owner, int field, int i, int j, double** value_dev) {
  const int epoch = ++c->bepoch;
  int* myflags = c->flags + c->L.ni;
  double* mymail = reinterpret_cast<double*>(c->flags + 2 * (size_t)c->L.ni + CFD_FLAG_SLOTS);
  if (owner) {
    if (c->bcast_n < 1) {
      cfd_set_error("cfd_scalar_bcast: the owner has no ranks attached (cfd_bcast_attach)");
      return CFD_ERR_INVALID;
    }
    if (field < 0 || field >= CFD_FIELD_COUNT || !c->field[field] || i < 1 || i > c->L.ni || j < 1 || j > c->L.nj) {
      cfd_set_error("cfd_scalar_bcast: bad field / index");
      return CFD_ERR_INVALID;
    }
    BcastParams B{};
    B.n = c->bcast_n;
    B.src = c->field[field] + (size_t)(i - 1) * c->L.LD + lpos(j - 1, c->L.M, c->L.T);
    B.epoch = epoch;
    for (int r = 0; r < c->bcast_n; ++r) {
      if (!c->bcast_flags[r]) {
        cfd_set_error("cfd_scalar_bcast: rank %d is not attached", r);
        return CFD_ERR_INVALID;
      }
      B.flags[r] = c->bcast_flags[r] + c->bcast_ni[r] + CFD_FLAG_BCAST;
      B.mail[r] = reinterpret_cast<double*>(c->bcast_flags[r] + 2 * (size_t)c->bcast_ni[r] + CFD_FLAG_SLOTS);
    }
    k_scalar_send<<<1, CFD_BCAST_MAX, 0, c->stream>>>(B);
    ++c->launches;
  }
  k_scalar_wait<<<1, 1, 0, c->stream>>>(myflags + CFD_FLAG_BCAST, epoch, myflags + CFD_FLAG_ABORT);
  ++c->launches;
  CFD_CUDA(cudaGetLastError());
  *value_dev = mymail + (epoch & 3);
  return CFD_OK;
}
