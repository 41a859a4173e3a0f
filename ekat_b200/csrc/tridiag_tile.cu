// tridiag_tile.cu -- tile-resident tridiagonal solvers: cyclic reduction (both layouts) and
// Thomas / bfb on the reference's column-major layout.
//
// A CTA stages a tile of C whole columns (all 3*nA diagonal planes and nrhs data planes)
// in shared memory as [plane][level][Cp] (column fastest, Cp odd => conflict-free both for
// the transposing fill and for per-column sweeps), solves in place on chip, and writes the
// solutions back. For the column-major layout the fill/drain are transposes with lanes
// along the contiguous level index, so drop-in callers holding [col][lev] views
// (tests/algorithm/tridiag_tests_performance.cpp:181-185) get coalesced HBM traffic
// without a separate transpose pass. The reference's team_barrier per CR level
// (ekat_tridiag.hpp:422,456) becomes one __syncthreads per level over the whole tile.
//
//  CR     (ekat_tridiag.hpp:385-630): W threads cooperate on each column (W = 1 when there
//         are many columns, up to the whole CTA when there are few), rows of a level are
//         strided over them exactly as over the reference's team; update formulas, operand
//         order and the clamped-neighbour trick (:408-418) are kept literally and unfused,
//         so results are bit-identical to the reference's CR at -ffp-contract=off.
//  BFB    (:292-320, :632-678): factorise (one thread per matrix), barrier, per-rhs solves.
//  THOMAS (:177-290): same structure, reciprocal + FMA form (tolerance-level agreement).
#include "common.cuh"

namespace ekb {

enum TileAlgo { ALGO_THOMAS = 0, ALGO_BFB = 1, ALGO_CR = 2 };

struct TileArgs {
  void *dl, *d, *du, *X;
  int64_t ncol, a_stride, x_stride;
  int nlev, nA, nrhs, C, Cp, lev_major;
  int clobber;  // planes written back under EKAT_B200_CLOBBER_DIAGS: bit 0 dl, bit 1 d, bit 2 du (0: none)
};

template <typename R>
__device__ __forceinline__ void tile_fill(const TileArgs& a, R* sdl, R* sd, R* sdu, R* sx, int64_t c0, int nc) {
  const R* gdl = static_cast<const R*>(a.dl);
  const R* gd = static_cast<const R*>(a.d);
  const R* gdu = static_cast<const R*>(a.du);
  const R* gx = static_cast<const R*>(a.X);
  const int nlev = a.nlev, nA = a.nA, nrhs = a.nrhs, Cp = a.Cp;
  const int tid = threadIdx.x, nthr = blockDim.x;
  if (a.lev_major) {
    // global [plane][lev][stride] -> smem [plane][lev][Cp]: lanes along columns
    const int na = nA * nlev, nx = nrhs * nlev;
    for (int e = tid; e < na * nc; e += nthr) {
      const int r = e / nc, c = e - r * nc;
      const size_t g = (size_t)r * a.a_stride + c0 + c;
      sdl[(size_t)r * Cp + c] = ld_stream(gdl + g);
      sd[(size_t)r * Cp + c] = ld_stream(gd + g);
      sdu[(size_t)r * Cp + c] = ld_stream(gdu + g);
    }
    for (int e = tid; e < nx * nc; e += nthr) {
      const int r = e / nc, c = e - r * nc;
      sx[(size_t)r * Cp + c] = ld_stream(gx + (size_t)r * a.x_stride + c0 + c);
    }
  } else {
    // global [col][lev][p] -> smem [p][lev][Cp]: lanes along the contiguous (lev,p) index
    const int na = nlev * nA, nx = nlev * nrhs;
    for (int e = tid; e < na * nc; e += nthr) {
      const int c = e / na, k = e - c * na;
      const int i = k / nA, p = k - i * nA;
      const size_t g = (size_t)(c0 + c) * a.a_stride + k;
      const size_t s = ((size_t)p * nlev + i) * Cp + c;
      sdl[s] = ld_stream(gdl + g);
      sd[s] = ld_stream(gd + g);
      sdu[s] = ld_stream(gdu + g);
    }
    for (int e = tid; e < nx * nc; e += nthr) {
      const int c = e / nx, k = e - c * nx;
      const int i = k / nrhs, j = k - i * nrhs;
      sx[((size_t)j * nlev + i) * Cp + c] = ld_stream(gx + (size_t)(c0 + c) * a.x_stride + k);
    }
  }
}

template <typename R>
__device__ __forceinline__ void tile_drain(const TileArgs& a, const R* sdl, const R* sd, const R* sdu, const R* sx,
                                           int64_t c0, int nc) {
  R* gdl = static_cast<R*>(a.dl);
  R* gd = static_cast<R*>(a.d);
  R* gdu = static_cast<R*>(a.du);
  R* gx = static_cast<R*>(a.X);
  const int nlev = a.nlev, nA = a.nA, nrhs = a.nrhs, Cp = a.Cp;
  const int tid = threadIdx.x, nthr = blockDim.x;
  if (a.lev_major) {
    const int na = nA * nlev, nx = nrhs * nlev;
    for (int e = tid; e < nx * nc; e += nthr) {
      const int r = e / nc, c = e - r * nc;
      st_stream(gx + (size_t)r * a.x_stride + c0 + c, sx[(size_t)r * Cp + c]);
    }
    if (a.clobber)
      for (int e = tid; e < na * nc; e += nthr) {
        const int r = e / nc, c = e - r * nc;
        const size_t g = (size_t)r * a.a_stride + c0 + c;
        if (a.clobber & 1) st_stream(gdl + g, sdl[(size_t)r * Cp + c]);
        if (a.clobber & 2) st_stream(gd + g, sd[(size_t)r * Cp + c]);
        if (a.clobber & 4) st_stream(gdu + g, sdu[(size_t)r * Cp + c]);
      }
  } else {
    const int na = nlev * nA, nx = nlev * nrhs;
    for (int e = tid; e < nx * nc; e += nthr) {
      const int c = e / nx, k = e - c * nx;
      const int i = k / nrhs, j = k - i * nrhs;
      st_stream(gx + (size_t)(c0 + c) * a.x_stride + k, sx[((size_t)j * nlev + i) * Cp + c]);
    }
    if (a.clobber)
      for (int e = tid; e < na * nc; e += nthr) {
        const int c = e / na, k = e - c * na;
        const int i = k / nA, p = k - i * nA;
        const size_t g = (size_t)(c0 + c) * a.a_stride + k;
        const size_t s = ((size_t)p * nlev + i) * Cp + c;
        if (a.clobber & 1) st_stream(gdl + g, sdl[s]);
        if (a.clobber & 2) st_stream(gd + g, sd[s]);
        if (a.clobber & 4) st_stream(gdu + g, sdu[s]);
      }
  }
}

// --------------------------------------------------------------------------- Thomas / bfb in smem
template <typename R, int ALGO>
__device__ __forceinline__ void tile_thomas(const TileArgs& a, R* sdl, R* sd, R* sdu, R* sx, int nc) {
  const int nlev = a.nlev, nA = a.nA, nrhs = a.nrhs, Cp = a.Cp;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const size_t L = Cp;  // stride between levels of one column
  // phase 1: factorise each matrix once (ekat_tridiag.hpp:177-193 / :292-303)
  for (int item = tid; item < nc * nA; item += nthr) {
    const int p = item / nc, c = item - p * nc;
    R* pl = sdl + (size_t)p * nlev * L + c;
    R* pd = sd + (size_t)p * nlev * L + c;
    R* pu = sdu + (size_t)p * nlev * L + c;
    if (ALGO == ALGO_BFB) {
      R dprev = pd[0];
      for (int i = 1; i < nlev; ++i) {
        const R q = x_div(pl[i * L], dprev);
        dprev = x_sub(pd[i * L], x_mul(q, pu[(i - 1) * L]));
        pl[i * L] = q;
        pd[i * L] = dprev;
      }
    } else {
      // dl <- m = dl*r, d <- r = 1/pivot, du <- c = du*r
      R r = f_rcp(pd[0]);
      R cprev = pu[0] * r;
      pd[0] = r;
      pu[0] = cprev;
      for (int i = 1; i < nlev; ++i) {
        const R l = pl[i * L];
        r = f_rcp(f_fma(-l, cprev, pd[i * L]));
        cprev = pu[i * L] * r;
        pl[i * L] = l * r;
        pd[i * L] = r;
        pu[i * L] = cprev;
      }
    }
  }
  __syncthreads();
  // phase 2: one (column, rhs) recurrence per thread (:195-214 / :305-320)
  for (int item = tid; item < nc * nrhs; item += nthr) {
    const int j = item / nc, c = item - j * nc;
    const int p = nA > 1 ? j : 0;
    const R* pl = sdl + (size_t)p * nlev * L + c;
    const R* pd = sd + (size_t)p * nlev * L + c;
    const R* pu = sdu + (size_t)p * nlev * L + c;
    R* px = sx + (size_t)j * nlev * L + c;
    if (ALGO == ALGO_BFB) {
      R xv = px[0];
      for (int i = 1; i < nlev; ++i) {
        xv = x_sub(px[i * L], x_mul(pl[i * L], xv));
        px[i * L] = xv;
      }
      xv = x_div(xv, pd[(nlev - 1) * L]);
      px[(nlev - 1) * L] = xv;
      for (int i = nlev - 1; i > 0; --i) {
        xv = x_div(x_sub(px[(i - 1) * L], x_mul(pu[(i - 1) * L], xv)), pd[(i - 1) * L]);
        px[(i - 1) * L] = xv;
      }
    } else {
      R xv = px[0] * pd[0];
      px[0] = xv;
      for (int i = 1; i < nlev; ++i) {
        xv = f_fma(-pl[i * L], xv, px[i * L] * pd[i * L]);
        px[i * L] = xv;
      }
      for (int i = nlev - 2; i >= 0; --i) {
        xv = f_fma(-pu[i * L], xv, px[i * L]);
        px[i * L] = xv;
      }
    }
  }
}

// --------------------------------------------------------------------------- cyclic reduction in smem
// W = blockDim.x / C threads per column; thread (c, w) takes rows stride*(w + W*m) of a level
// (the reference's `for (i = stride*team_id; i < nrow; i += stride*nteam)`, :403, :483, :572).
template <typename R>
__device__ __forceinline__ void tile_cr(const TileArgs& a, R* sdl, R* sd, R* sdu, R* sx, int nc) {
  const int nrow = a.nlev, nA = a.nA, nrhs = a.nrhs, Cp = a.Cp, C = a.C;
  const int W = blockDim.x / C;
  const int c = threadIdx.x % C, w = threadIdx.x / C;
  const bool act = c < nc && w < W;
  const size_t L = Cp, PA = (size_t)nrow * L;  // level stride, plane stride
  R* pl = sdl + c;
  R* pd = sd + c;
  R* pu = sdu + c;
  R* px = sx + c;
  int os = 1, stride;
  while ((stride = (os << 1)) < nrow) {
    if (act) {
      for (int i = stride * w; i < nrow; i += stride * W) {
        const bool im_ok = i - os >= 0, ip_ok = i + os < nrow;
        const int im = im_ok ? i - os : i, ip = ip_ok ? i + os : i;  // :413-414
        if (nA == 1) {
          const R f1 = im_ok ? x_div(-pl[i * L], pd[im * L]) : (R)0;  // :408-409
          const R f2 = ip_ok ? x_div(-pu[i * L], pd[ip * L]) : (R)0;
          for (int j = 0; j < nrhs; ++j) {
            R* xj = px + j * PA;
            xj[i * L] = x_add(xj[i * L], x_add(x_mul(f1, xj[im * L]), x_mul(f2, xj[ip * L])));  // :418 / :495
          }
          pl[i * L] = x_mul(f1, pl[im * L]);  // :415-417 / :500-502, statement order kept
          pu[i * L] = x_mul(f2, pu[ip * L]);
          pd[i * L] = x_add(pd[i * L], x_add(x_mul(f1, pu[im * L]), x_mul(f2, pl[ip * L])));
        } else {
          for (int j = 0; j < nrhs; ++j) {  // :579-586
            R* lj = pl + j * PA;
            R* dj = pd + j * PA;
            R* uj = pu + j * PA;
            R* xj = px + j * PA;
            const R f1 = im_ok ? x_div(-lj[i * L], dj[im * L]) : (R)0;
            const R f2 = ip_ok ? x_div(-uj[i * L], dj[ip * L]) : (R)0;
            lj[i * L] = x_mul(f1, lj[im * L]);
            uj[i * L] = x_mul(f2, uj[ip * L]);
            dj[i * L] = x_add(dj[i * L], x_add(x_mul(f1, uj[im * L]), x_mul(f2, lj[ip * L])));
            xj[i * L] = x_add(xj[i * L], x_add(x_mul(f1, xj[im * L]), x_mul(f2, xj[ip * L])));
          }
        }
      }
    }
    os <<= 1;
    __syncthreads();  // team_barrier :422
  }
  // bottom 1x1 or 2x2 solve (:427-437 / :508-521 / :592-605), rhs strided over the column's threads
  if (act) {
    for (int j = w; j < nrhs; j += W) {
      const size_t ap = (nA > 1 ? j : 0) * PA;
      R* xj = px + j * PA;
      if (os >= nrow) {
        xj[0] = x_div(xj[0], pd[ap]);
      } else {
        const R d0 = pd[ap], dos = pd[ap + os * L], du0 = pu[ap], dlos = pl[ap + os * L];
        const R det = x_sub(x_mul(d0, dos), x_mul(du0, dlos));
        const R x0 = xj[0], x1 = xj[os * L];
        xj[0] = x_div(x_sub(x_mul(dos, x0), x_mul(du0, x1)), det);
        xj[os * L] = x_div(x_sub(x_mul(d0, x1), x_mul(dlos, x0)), det);
      }
    }
  }
  __syncthreads();  // :438
  os >>= 1;
  while (os) {  // up-sweep :442-457
    stride = os << 1;
    if (act) {
      for (int i = stride * w + os; i < nrow; i += stride * W) {
        const int im = i - os, ip = i + os;
        const bool im_ok = im >= 0, ip_ok = ip < nrow;
        for (int j = 0; j < nrhs; ++j) {
          const size_t ap = (nA > 1 ? j : 0) * PA;
          R* xj = px + j * PA;
          R f = 0;
          f = x_add(f, im_ok ? x_mul(pl[ap + i * L], xj[im * L]) : (R)0);  // :450-451
          f = x_add(f, ip_ok ? x_mul(pu[ap + i * L], xj[ip * L]) : (R)0);
          xj[i * L] = x_div(x_sub(xj[i * L], f), pd[ap + i * L]);  // :452
        }
      }
    }
    os >>= 1;
    __syncthreads();  // :456
  }
}

template <typename R, int ALGO>
__global__ void __launch_bounds__(256) k_tridiag_tile(const TileArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const size_t plane = (size_t)a.nlev * a.Cp;
  R* sdl = reinterpret_cast<R*>(smem_raw);
  R* sd = sdl + plane * a.nA;
  R* sdu = sd + plane * a.nA;
  R* sx = sdu + plane * a.nA;
  const int64_t ntiles = (a.ncol + a.C - 1) / a.C;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t c0 = tile * a.C;
    const int nc = (int)tmin<int64_t>(a.C, a.ncol - c0);
    tile_fill<R>(a, sdl, sd, sdu, sx, c0, nc);
    __syncthreads();
    if (ALGO == ALGO_CR) tile_cr<R>(a, sdl, sd, sdu, sx, nc);
    else tile_thomas<R, ALGO>(a, sdl, sd, sdu, sx, nc);
    __syncthreads();
    tile_drain<R>(a, sdl, sd, sdu, sx, c0, nc);
    __syncthreads();
  }
}

template <typename R>
int tridiag_tile(int algo, R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int nA, int nrhs, int64_t a_stride,
                 int64_t x_stride, int layout, unsigned flags, cudaStream_t stream) {
  const size_t per_col = (size_t)(3 * nA + nrhs) * nlev * sizeof(R);
  const Tuning tu = tuning();
  // Columns per tile. Many columns: up to a warp of them, and small enough that >= 2 CTAs share an SM.
  // Few columns: spread them over the SMs and give each column many threads (CR) instead.
  if (per_col > kMaxSmemPerCta) return EKAT_B200_ERR_TOO_LARGE;
  int C = 32;
  while (C > 1 && (size_t)2 * (C | 1) * per_col > kMaxSmemPerCta) C /= 2;  // aim for >= 2 CTAs per SM
  const int64_t spread = (ncol + kNumSms - 1) / kNumSms;
  if (spread < C) C = (int)tmax<int64_t>(1, spread);
  if (tu.cols_per_cta > 0) C = tu.cols_per_cta;
  int Cp = C > 1 ? (C | 1) : 1;
  while (C > 1 && (size_t)Cp * per_col > kMaxSmemPerCta) {
    --C;
    Cp = C > 1 ? (C | 1) : 1;
  }
  const size_t smem = (size_t)Cp * per_col;

  TileArgs a;
  a.dl = dl; a.d = d; a.du = du; a.X = X;
  a.ncol = ncol; a.a_stride = a_stride; a.x_stride = x_stride;
  a.nlev = nlev; a.nA = nA; a.nrhs = nrhs; a.C = C; a.Cp = Cp;
  a.lev_major = layout == EKAT_B200_LAYOUT_LEV_MAJOR;
  // What the reference's in-place solvers leave behind (ekat_tridiag.hpp): serial thomas -- pivots in d (:216-227);
  // bfb / team thomas -- multipliers in dl, pivots in d (:292-303, :177-193); cr -- all three reduced (:415-417).
  // du is never modified by thomas / bfb. The reciprocal-form sweep below keeps 1/pivot, m*r and du*r in the staged
  // planes, so a thomas call that must write the pivots back runs the bfb sweep instead (same operations, each
  // rounded once: within the thomas tolerance by construction, and d comes out bit-identical to the reference's).
  const bool clobber = (flags & EKAT_B200_CLOBBER_DIAGS) != 0;
  a.clobber = !clobber ? 0 : (algo == ALGO_CR ? 7 : (algo == ALGO_BFB ? 3 : 2));
  if (clobber && algo == ALGO_THOMAS) algo = ALGO_BFB;

  const int64_t ntiles = (ncol + C - 1) / C;
  int per_sm = (int)(kMaxSmemPerCta / (smem + 1024));
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 8) per_sm = 8;
  const unsigned grid = (unsigned)tmin<int64_t>(ntiles, (int64_t)kNumSms * per_sm);
  int rc;
  const char* name;
  const bool f64 = sizeof(R) == 8;
  if (algo == ALGO_CR) {
    auto k = k_tridiag_tile<R, ALGO_CR>;
    if ((rc = allow_smem(k, smem))) return rc;
    k<<<grid, 256, smem, stream>>>(a);
    name = f64 ? "tridiag_tile<f64,cr>" : "tridiag_tile<f32,cr>";
  } else if (algo == ALGO_BFB) {
    auto k = k_tridiag_tile<R, ALGO_BFB>;
    if ((rc = allow_smem(k, smem))) return rc;
    k<<<grid, 256, smem, stream>>>(a);
    name = f64 ? "tridiag_tile<f64,bfb>" : "tridiag_tile<f32,bfb>";
  } else {
    auto k = k_tridiag_tile<R, ALGO_THOMAS>;
    if ((rc = allow_smem(k, smem))) return rc;
    k<<<grid, 256, smem, stream>>>(a);
    name = f64 ? "tridiag_tile<f64,thomas>" : "tridiag_tile<f32,thomas>";
  }
  note_launch(name);
  return check_cuda(cudaGetLastError());
}

template int tridiag_tile<double>(int, double*, double*, double*, double*, int64_t, int, int, int, int64_t, int64_t,
                                  int, unsigned, cudaStream_t);
template int tridiag_tile<float>(int, float*, float*, float*, float*, int64_t, int, int, int, int64_t, int64_t, int,
                                 unsigned, cudaStream_t);

}  // namespace ekb
