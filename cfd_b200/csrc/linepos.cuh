// linepos.cuh -- position of node j0 inside a line (line layout, common.cuh) for a thread that owns the chunk
// starting at jb = t*M: nodes of the own chunk and of the two neighbouring chunks are reached with adds and
// selects; anything farther away (only possible on tiny grids, M < 3) falls back to the div/mod form.
#pragma once

namespace cfd {

struct LinePos {
  int M, T, t, jb, wrap;   // wrap = M*T - 1
  __device__ __forceinline__ LinePos(int M_, int T_, int t_) : M(M_), T(T_), t(t_), jb(t_ * M_), wrap(M_ * T_ - 1) {}
  __device__ __forceinline__ int operator()(int j0) const {
    const int d = j0 - jb;
    int pos = d * T + t;
    if (d < 0) pos += wrap;
    if (d >= M) pos -= wrap;
    if ((unsigned)(d + M) >= (unsigned)(3 * M)) pos = (j0 % M) * T + (j0 / M);
    return pos;
  }
};

}  // namespace cfd
