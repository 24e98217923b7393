// Shared device helpers for the D4 engine (sm_100a, FP64).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define D4_MREF 7
#define D4_WARP 32
#define D4_FULL 0xffffffffu

namespace d4k {

// ---- exact (non-contracted) arithmetic for the cutoff tests: the reference evaluates
//      vec = xi - (xj + t); r2 = vec1*vec1 + vec2*vec2 + vec3*vec3 without FMA (rational.f90:290-292) -------------
__device__ __forceinline__ double exact_r2(double vx, double vy, double vz) {
  return __dadd_rn(__dadd_rn(__dmul_rn(vx, vx), __dmul_rn(vy, vy)), __dmul_rn(vz, vz));
}

// A lattice image of atom j: position = fl(xyz[j] + trans[t]) exactly as the reference forms it.
// w packs (j: 32 bit | species: 8 bit | t: 24 bit).
struct __align__(32) Point {
  double x, y, z;
  unsigned long long w;
};
__device__ __forceinline__ unsigned long long pack_point(int j, int species, int t) {
  return (unsigned long long)(unsigned)j | ((unsigned long long)(species & 0xff) << 32) |
         ((unsigned long long)(t & 0xffffff) << 40);
}
__device__ __forceinline__ int point_atom(unsigned long long w) { return (int)(w & 0xffffffffull); }
__device__ __forceinline__ int point_species(unsigned long long w) { return (int)((w >> 32) & 0xff); }
__device__ __forceinline__ int point_trans(unsigned long long w) { return (int)((w >> 40) & 0xffffff); }

__device__ __forceinline__ Point load_point(const Point* p) {
  // one 32-byte vector load
  const double4 v = *reinterpret_cast<const double4*>(p);
  Point r;
  r.x = v.x; r.y = v.y; r.z = v.z;
  r.w = (unsigned long long)__double_as_longlong(v.w);
  return r;
}

// Frame-exact squared distance of the directed entry (row atom i -> image point p of atom j).  The reference only
// walks pairs with jat <= iat and forms vec = xyz(iat) - (xyz(jat) + trans) (rational.f90:290, atm.f90:332), i.e. in
// the frame of the HIGHER-index atom.  For j > i the mirrored entry is what the reference evaluates:
// vec = x_j - fl(x_i - T_t) (the translation list holds -T exactly).  The two roundings can only disagree about the
// cutoff test when r2 is within a few ulp of cutoff^2, so the mirrored form is evaluated in that window only.
__device__ __forceinline__ double frame_r2(double r2, double cut2, int i, double xi, double yi, double zi, const Point& p,
                                           const double* __restrict__ xyz, const double* __restrict__ trans) {
  if (fabs(r2 - cut2) <= cut2 * 1.0e-10) {
    const int j = point_atom(p.w);
    if (j > i) {
      const int t = point_trans(p.w);
      r2 = exact_r2(__dsub_rn(xyz[3 * j], __dsub_rn(xi, trans[3 * t])), __dsub_rn(xyz[3 * j + 1], __dsub_rn(yi, trans[3 * t + 1])),
                    __dsub_rn(xyz[3 * j + 2], __dsub_rn(zi, trans[3 * t + 2])));
    }
  }
  return r2;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(D4_FULL, v, o);
  return v;
}
__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(D4_FULL, v, o);
  return v;
}

// Quintic smooth switch, src/dftd4/cutoff.f90:74-114
__device__ __forceinline__ void smooth_cutoff(double r, double cutoff, double width, double& sw, double& dsw) {
  if (width <= 0.0 || cutoff <= 0.0) { sw = 1.0; dsw = 0.0; return; }
  const double w = fmin(width, cutoff);
  const double inner = cutoff - w;
  if (r <= inner) { sw = 1.0; dsw = 0.0; }
  else if (r >= cutoff) { sw = 0.0; dsw = 0.0; }
  else {
    const double x = (cutoff - r) / w;
    sw = x * x * x * (10.0 + x * (-15.0 + 6.0 * x));
    dsw = -30.0 * x * x * (1.0 - x) * (1.0 - x) / w;
  }
}

// Uniform grid over the image-expanded point set
struct Grid {
  double ox, oy, oz;      // origin
  double h, inv_h;        // cell edge
  int nx, ny, nz;         // cells per dimension
  double lox, loy, loz, hix, hiy, hiz;  // acceptance box for image points
  int reach;              // cells to visit on each side
};
__device__ __forceinline__ int grid_coord(double x, double o, double inv_h, int n) {
  int c = (int)floor((x - o) * inv_h);
  return c < 0 ? 0 : (c >= n ? n - 1 : c);
}

}  // namespace d4k
