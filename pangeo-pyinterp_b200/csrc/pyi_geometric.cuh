// pyi_geometric.cuh -- K1-K3: geometric interpolation of one query point
// (pybind/geometric/bivariate.hpp:47-97, trivariate.hpp:47-123,
// quadrivariate.hpp:50-148 of the reference), split into its three stages so
// that the kernels can schedule the corner gather on their own terms:
//
//   geo_lookup    axis brackets (Axis::find_indexes per axis)
//   geo_gather_*  the 2^ND corner values, from the caller's C-order array, from
//                 the cell-packed copy by one thread, or from the cell-packed
//                 copy by a group of lanes (one line request per cell)
//   geo_evaluate  corner coordinates, casts to the result type and the
//                 evaluators of math/interpolate/geometric/*.hpp
//
// Included by pyi_kernels.cu (compiled with -fmad=false).
#pragma once

namespace pyi {

template <bool ZI64>
struct ZOf { using type = double; };
template <>
struct ZOf<true> { using type = long long; };

struct Brackets {
  int ix[2], iy[2], iz[2], iu[2];
};

// Returns -1, or the first axis (x, y, z, u order) that cannot bracket the point.
template <int ND, bool ZI64>
__device__ __forceinline__ int geo_lookup(const GeoParams& p, double x, double y,
                                          typename ZOf<ZI64>::type z, double u, Brackets& b) {
  using Z = typename ZOf<ZI64>::type;
  b.iz[0] = b.iz[1] = b.iu[0] = b.iu[1] = 0;
  if (!find_indexes<double>(p.ax[0], x, b.ix[0], b.ix[1])) return 0;
  if (!find_indexes<double>(p.ax[1], y, b.iy[0], b.iy[1])) return 1;
  if (ND >= 3) {
    if (!find_indexes<Z>(p.ax[2], z, b.iz[0], b.iz[1])) return 2;
  }
  if (ND >= 4) {
    if (!find_indexes<double>(p.ax[3], u, b.iu[0], b.iu[1])) return 3;
  }
  return -1;
}

// Corner block q[c][d][a][b]: a = x, b = y, c = z, d = u corner.
template <int ND, typename R>
struct Corners {
  static constexpr int NZ = ND >= 3 ? 2 : 1;
  static constexpr int NU = ND >= 4 ? 2 : 1;
  static constexpr int NV = 4 * NZ * NU;
  static constexpr int BYTES = NV * static_cast<int>(sizeof(R));
  R q[NZ][NU][2][2];
};

// From the caller's C-order array (Grid::value, pybind/grid.hpp:340-345).
template <int ND, int DT, typename R>
__device__ __forceinline__ void geo_gather_plain(const GeoParams& p, const Brackets& b,
                                                 Corners<ND, R>& k) {
  const long long n1 = p.ax[1].n;
  const long long n2 = ND >= 3 ? p.ax[2].n : 1;
  const long long n3 = ND >= 4 ? p.ax[3].n : 1;
#pragma unroll
  for (int c = 0; c < Corners<ND, R>::NZ; ++c) {
#pragma unroll
    for (int d = 0; d < Corners<ND, R>::NU; ++d) {
#pragma unroll
      for (int a = 0; a < 2; ++a) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          long long off = b.ix[a] * n1 + b.iy[e];
          if (ND >= 3) off = off * n2 + b.iz[c];
          if (ND >= 4) off = off * n3 + b.iu[d];
          k.q[c][d][a][e] = load_value<DT, R>(p.data, p.dtype, off);
        }
      }
    }
  }
}

// Cell index and orientation of a bracket pair on one axis for the cell-packed
// copy: a pair is either (k, k+1) / the periodic seam (n-1, 0) -> cell k, or
// the same pair reversed (descending axes, axis.hpp:759-777) -> cell of the
// lower index with the two corners exchanged.
__device__ __forceinline__ int cell_of(const DAxis& a, int i0, int i1, bool& swapped) {
  const bool fwd = (i1 == i0 + 1) || (a.periodic && i0 == a.n - 1 && i1 == 0);
  swapped = !fwd;
  return fwd ? i0 : i1;
}

struct CellRef {
  long long cell;
  bool sx, sy, sz, su;
};

template <int ND>
__device__ __forceinline__ CellRef cell_ref(const GeoParams& p, const Brackets& b) {
  CellRef r;
  r.sz = r.su = false;
  r.cell = static_cast<long long>(cell_of(p.ax[0], b.ix[0], b.ix[1], r.sx)) * p.ax[1].n +
           cell_of(p.ax[1], b.iy[0], b.iy[1], r.sy);
  if (ND >= 3) r.cell = r.cell * p.ax[2].n + cell_of(p.ax[2], b.iz[0], b.iz[1], r.sz);
  if (ND >= 4) r.cell = r.cell * p.ax[3].n + cell_of(p.ax[3], b.iu[0], b.iu[1], r.su);
  return r;
}

// One 256-bit load.  The memory system serves a fixed number of load REQUESTS
// per second for scattered lines (~36 G/s on a B200, independent of the bytes
// asked for: tools/gather_probe.cu, profiles/r1_gather_probe.txt), so 32 bytes
// are read with one request, not two; .L2::64B keeps the DRAM fill at 64 bytes
// instead of the whole 128-byte line.
__device__ __forceinline__ void load_256(const void* src, unsigned long long* w) {
  asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v4.u64 {%0,%1,%2,%3}, [%4];"
               : "=l"(w[0]), "=l"(w[1]), "=l"(w[2]), "=l"(w[3])
               : "l"(src));
}

// number of cells of the packed copy: one per grid node (the last one along an
// axis is the seam cell of a periodic axis, unused otherwise)
template <int ND>
__device__ __forceinline__ long long geo_total_cells(const GeoParams& p) {
  long long t = static_cast<long long>(p.ax[0].n) * p.ax[1].n;
  if (ND >= 3) t *= p.ax[2].n;
  if (ND >= 4) t *= p.ax[3].n;
  return t;
}

// Packed value block -> q, exchanging corners of reversed pairs.
template <int ND, typename R>
__device__ __forceinline__ void unpack_cell(const R* v, const CellRef& r, Corners<ND, R>& k) {
  constexpr int NZ = Corners<ND, R>::NZ;
  constexpr int NU = Corners<ND, R>::NU;
#pragma unroll
  for (int a = 0; a < 2; ++a) {
#pragma unroll
    for (int e = 0; e < 2; ++e) {
#pragma unroll
      for (int c = 0; c < NZ; ++c) {
#pragma unroll
        for (int d = 0; d < NU; ++d) k.q[c][d][a][e] = v[((a * 2 + e) * NZ + c) * NU + d];
      }
    }
  }
  if (r.sx | r.sy | r.sz | r.su) {
#pragma unroll
    for (int c = 0; c < NZ; ++c) {
#pragma unroll
      for (int d = 0; d < NU; ++d) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const R t0 = k.q[c][d][0][e], t1 = k.q[c][d][1][e];
          k.q[c][d][0][e] = r.sx ? t1 : t0;
          k.q[c][d][1][e] = r.sx ? t0 : t1;
        }
#pragma unroll
        for (int a = 0; a < 2; ++a) {
          const R t0 = k.q[c][d][a][0], t1 = k.q[c][d][a][1];
          k.q[c][d][a][0] = r.sy ? t1 : t0;
          k.q[c][d][a][1] = r.sy ? t0 : t1;
        }
      }
    }
    if (ND >= 3) {
#pragma unroll
      for (int d = 0; d < NU; ++d) {
#pragma unroll
        for (int a = 0; a < 2; ++a) {
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const R t0 = k.q[0][d][a][e], t1 = k.q[NZ - 1][d][a][e];
            k.q[0][d][a][e] = r.sz ? t1 : t0;
            k.q[NZ - 1][d][a][e] = r.sz ? t0 : t1;
          }
        }
      }
    }
    if (ND >= 4) {
#pragma unroll
      for (int c = 0; c < NZ; ++c) {
#pragma unroll
        for (int a = 0; a < 2; ++a) {
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const R t0 = k.q[c][0][a][e], t1 = k.q[c][NU - 1][a][e];
            k.q[c][0][a][e] = r.su ? t1 : t0;
            k.q[c][NU - 1][a][e] = r.su ? t0 : t1;
          }
        }
      }
    }
  }
}

// From the cell-packed copy, one thread per cell.
template <int ND, typename R>
__device__ __forceinline__ void geo_gather_packed(const GeoParams& p, const CellRef& r,
                                                  Corners<ND, R>& k) {
  constexpr int NV = Corners<ND, R>::NV;
  constexpr int BYTES = Corners<ND, R>::BYTES;
  union {
    unsigned long long w[BYTES / 8];
    R v[NV];
  } buf;
  PYI_CHECK_RANGE(r.cell, 1, geo_total_cells<ND>(p), "packed cell (one thread)");
  const char* src = reinterpret_cast<const char*>(static_cast<const R*>(p.packed) + r.cell * NV);
  if (BYTES >= 32) {
#pragma unroll
    for (int c = 0; c < BYTES / 32; ++c) load_256(src + 32 * c, buf.w + 4 * c);
  } else {  // 2-D float cells: 16 bytes
    asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0,%1}, [%2];"
                 : "=l"(buf.w[0]), "=l"(buf.w[1])
                 : "l"(src));
  }
  unpack_cell<ND, R>(buf.v, r, k);
}

// From the cell-packed copy, G = BYTES / 32 adjacent lanes per cell.  A cell of
// 64 (128) bytes read by one thread is 2 (4) line requests; read by a lane
// pair (quad) in ONE instruction -- each lane a different 32-byte chunk of the
// SAME cell -- it is one request, which doubles (quadruples) the gather rate
// (profiles/r1_gather_probe.txt: 18 -> 35 Ggather/s for 64-byte cells).
// Round g serves the cell of group member g; the chunks are then exchanged
// with shuffles.  Must be called by all 32 lanes of the warp (a lane without a
// valid point passes cell 0).
template <int ND, typename R>
__device__ __forceinline__ void geo_gather_group(const GeoParams& p, const CellRef& r,
                                                 Corners<ND, R>& k) {
  constexpr int NV = Corners<ND, R>::NV;
  constexpr int BYTES = Corners<ND, R>::BYTES;
  constexpr int G = BYTES / 32;
  static_assert(G == 2 || G == 4, "group gather is for 64- and 128-byte cells");
  const unsigned lane = threadIdx.x & 31u;
  const unsigned m = lane & (G - 1);  // my chunk / member index
  const unsigned base = lane & ~static_cast<unsigned>(G - 1);
  unsigned long long tmp[G][4];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const long long c = __shfl_sync(0xffffffffu, r.cell, base + g);
    PYI_CHECK_RANGE(c, 1, geo_total_cells<ND>(p), "packed cell (lane group)");
    load_256(reinterpret_cast<const char*>(p.packed) + c * BYTES + m * 32, tmp[g]);
  }
  union {
    unsigned long long w[BYTES / 8];
    R v[NV];
  } buf;
  // After round g, lane (base + j) holds chunk j of member g's cell in tmp[g],
  // so chunk t of MY cell is tmp[m] of lane base + t.  Exchange with lane ^ d
  // (member m ^ d): it needs chunk m of its cell = my tmp[m ^ d]; it sends me
  // its tmp[m] = chunk m ^ d of my cell.  d = 0 keeps my own chunk m.
#pragma unroll
  for (int q = 0; q < BYTES / 8; ++q) buf.w[q] = 0;
#pragma unroll
  for (int d = 0; d < G; ++d) {
    const unsigned t = m ^ d;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      unsigned long long send = tmp[0][q];
#pragma unroll
      for (int g = 1; g < G; ++g) send = (t == static_cast<unsigned>(g)) ? tmp[g][q] : send;
      const unsigned long long got = d == 0 ? send : __shfl_xor_sync(0xffffffffu, send, d);
#pragma unroll
      for (int c = 0; c < G; ++c) buf.w[4 * c + q] = (t == static_cast<unsigned>(c)) ? got : buf.w[4 * c + q];
    }
  }
  unpack_cell<ND, R>(buf.v, r, k);
}

// Corner coordinates, casts and evaluators: bivariate.hpp:78-96,
// trivariate.hpp:90-121, quadrivariate.hpp:110-146 and
// math/interpolate/geometric/multivariate.hpp:133-153,271-306.
template <int ND, typename R, bool ZI64>
__device__ __forceinline__ R geo_evaluate(const GeoParams& p, const Brackets& b, double x, double y,
                                          typename ZOf<ZI64>::type z, double u,
                                          const Corners<ND, R>& k) {
  using Z = typename ZOf<ZI64>::type;
  constexpr int NZ = Corners<ND, R>::NZ;
  constexpr int NU = Corners<ND, R>::NU;
  // corner coordinates in R; x1 and the query x are normalised against the
  // R-rounded x0 promoted back to double (bivariate.hpp:78-89)
  const R x0 = static_cast<R>(axis_coord_f64(p.ax[0], b.ix[0]));
  const R x1 = static_cast<R>(
      axis_normalize(p.ax[0], axis_coord_f64(p.ax[0], b.ix[1]), static_cast<double>(x0)));
  const R y0 = static_cast<R>(axis_coord_f64(p.ax[1], b.iy[0]));
  const R y1 = static_cast<R>(axis_coord_f64(p.ax[1], b.iy[1]));
  const R px = static_cast<R>(axis_normalize(p.ax[0], x, static_cast<double>(x0)));
  const R py = static_cast<R>(y);
  R sp[NZ][NU];
#pragma unroll
  for (int c = 0; c < NZ; ++c) {
#pragma unroll
    for (int d = 0; d < NU; ++d) {
      sp[c][d] = eval_spatial<R>(p.method, p.idw_exp, px, py, x0, y0, x1, y1, k.q[c][d][0][0],
                                 k.q[c][d][0][1], k.q[c][d][1][0], k.q[c][d][1][1]);
    }
  }
  if (ND == 2) return sp[0][0];
  const Z z0 = axis_coord(p.ax[2], b.iz[0], Z{});
  const Z z1 = axis_coord(p.ax[2], b.iz[1], Z{});
  R along[NU];
#pragma unroll
  for (int d = 0; d < NU; ++d) {
    along[d] = axis_interp<Z, R>(p.z_method, z, z0, z1, sp[0][d], sp[NZ - 1][d]);
  }
  if (ND == 3) return along[0];
  const double u0 = axis_coord_f64(p.ax[3], b.iu[0]);
  const double u1 = axis_coord_f64(p.ax[3], b.iu[1]);
  return axis_interp<double, R>(p.u_method, u, u0, u1, along[0], along[NU - 1]);
}

// One query point, corners read by this thread alone.  Returns -1 and the
// value, or the first axis that could not bracket the point.
template <int ND, int DT, typename R, bool ZI64, bool PACKED = false>
__device__ __forceinline__ int geometric_eval(const GeoParams& p, double x, double y,
                                              typename ZOf<ZI64>::type z, double u, R& res) {
  Brackets b;
  const int bad = geo_lookup<ND, ZI64>(p, x, y, z, u, b);
  if (bad >= 0) return bad;
  Corners<ND, R> k;
  if (PACKED) {
    geo_gather_packed<ND, R>(p, cell_ref<ND>(p, b), k);
  } else {
    geo_gather_plain<ND, DT, R>(p, b, k);
  }
  res = geo_evaluate<ND, R, ZI64>(p, b, x, y, z, u, k);
  return -1;
}

// Generic kernel: one point per thread and iteration.
template <int ND, int DT, typename R, bool ZI64, bool PACKED>
__global__ void __launch_bounds__(kThreads)
geometric_kernel(const __grid_constant__ GeoParams p) {
  using Z = typename ZOf<ZI64>::type;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  unsigned long long key = kNoOob;
  R* __restrict__ out = static_cast<R*>(p.out);
  for (long long k = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; k < p.n;
       k += stride) {
    R res = RT<R>::nan();
    const Z z = ND >= 3 ? __ldg(static_cast<const Z*>(p.z) + k) : Z(0);
    const double u = ND >= 4 ? __ldg(p.u + k) : 0.0;
    const int bad =
        geometric_eval<ND, DT, R, ZI64, PACKED>(p, __ldg(p.x + k), __ldg(p.y + k), z, u, res);
    if (bad >= 0) {
      res = RT<R>::nan();
      const unsigned long long kk = (static_cast<unsigned long long>(p.index_base + k) << 2) | bad;
      key = kk < key ? kk : key;
    }
    __stcs(out + k, res);
  }
  if (p.bounds_error && key != kNoOob) {
    atomicMin(p.oob_key, key);
  }
}

// Packed cells of 64 / 128 bytes: whole warps iterate together so that lane
// groups can share the line requests (geo_gather_group).
template <int ND, typename R, bool ZI64>
__global__ void __launch_bounds__(kThreads)
geometric_group_kernel(const __grid_constant__ GeoParams p) {
  using Z = typename ZOf<ZI64>::type;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  unsigned long long key = kNoOob;
  R* __restrict__ out = static_cast<R*>(p.out);
  // the loop bound is warp-uniform: k - lane is the same for the 32 lanes
  const unsigned lane = threadIdx.x & 31u;
  for (long long k = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; k - lane < p.n;
       k += stride) {
    const bool active = k < p.n;
    const long long kk = active ? k : 0;
    const double x = __ldg(p.x + kk);
    const double y = __ldg(p.y + kk);
    const Z z = ND >= 3 ? __ldg(static_cast<const Z*>(p.z) + kk) : Z(0);
    const double u = ND >= 4 ? __ldg(p.u + kk) : 0.0;
    Brackets b;
    const int bad = geo_lookup<ND, ZI64>(p, x, y, z, u, b);
    CellRef r = cell_ref<ND>(p, b);
    if (bad >= 0) r.cell = 0;  // brackets are meaningless: any legal address
    Corners<ND, R> c;
    geo_gather_group<ND, R>(p, r, c);
    R res = RT<R>::nan();
    if (bad < 0) {
      res = geo_evaluate<ND, R, ZI64>(p, b, x, y, z, u, c);
    } else if (active) {
      const unsigned long long e = (static_cast<unsigned long long>(p.index_base + k) << 2) | bad;
      key = e < key ? e : key;
    }
    if (active) __stcs(out + k, res);
  }
  if (p.bounds_error && key != kNoOob) {
    atomicMin(p.oob_key, key);
  }
}

}  // namespace pyi
