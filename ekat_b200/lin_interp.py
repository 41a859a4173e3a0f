"""LinInterp on torch CUDA tensors -- the Python face of ekat::LinInterp
(reference: src/algorithm/ekat_lin_interp.hpp:38-157, ekat_lin_interp_impl.hpp).

    li = LinInterp(ncol, km1, km2, dtype, pack_size)     # allocates the index map, like the reference ctor
    li.setup(x1, x2)                                      # bracket search for every column -> li.indx_map
    li.lin_interp(x1, x2, y1, y2)                         # any number of fields with the same coordinates
    li.run(x1, x2, y1, y2)                                # fused setup + lin_interp (map stays on chip)

Views are the reference's scalarized Pack views: x1,y1 [ncol, km1_pack*N], x2,y2 [ncol, km2_pack*N].
y1/y2 may carry a leading field dimension [nfields, ncol, pitch].
"""
import torch

from ._lib import lib, check

_SFX = {torch.float64: "f64", torch.float32: "f32"}


def npack(n, pack_size):
    """ekat::npack (src/pack/ekat_pack.hpp:611-615)."""
    return (n + pack_size - 1) // pack_size


def _stream():
    return torch.cuda.current_stream().cuda_stream


class LinInterp:
    def __init__(self, ncol, km1, km2, dtype=torch.float64, pack_size=1, device="cuda"):
        if pack_size < 1 or pack_size & (pack_size - 1):
            raise ValueError("pack size must be a power of two (src/pack/ekat_pack.hpp:635)")
        if dtype not in _SFX:
            raise TypeError("float32 / float64 only")
        self.ncol, self.km1, self.km2 = ncol, km1, km2
        self.dtype, self.pack_size, self.device = dtype, pack_size, device
        self._km1_pack, self._km2_pack = npack(km1, pack_size), npack(km2, pack_size)
        lib()  # fail now, loudly, if the CUDA library is missing
        # m_indx_map (ekat_lin_interp.hpp:156); ints are kept unpadded: [ncol, km2]
        self.indx_map = torch.zeros((ncol, km2), dtype=torch.int32, device=device)

    def km1_pack(self):
        return self._km1_pack

    def km2_pack(self):
        return self._km2_pack

    @property
    def pitch1(self):
        return self._km1_pack * self.pack_size

    @property
    def pitch2(self):
        return self._km2_pack * self.pack_size

    # -- checks (the reference's EKAT_KERNEL_REQUIRE on pack divisibility, ekat_pack_kokkos.hpp:266-267) --
    def _coord(self, t, pitch, name):
        if t.dtype != self.dtype or not t.is_cuda:
            raise ValueError(f"{name}: expected a {self.dtype} CUDA tensor")
        if t.shape != (self.ncol, pitch) or (self.ncol and pitch and not t.is_contiguous()):
            raise ValueError(f"{name}: expected contiguous [{self.ncol}, {pitch}] (= npack*N), got {tuple(t.shape)}")

    def _field(self, t, pitch, name):
        if t.dtype != self.dtype or not t.is_cuda:
            raise ValueError(f"{name}: expected a {self.dtype} CUDA tensor")
        t3 = t if t.dim() == 3 else t.unsqueeze(0)
        if t3.shape[1:] != (self.ncol, pitch) or (t3.numel() and not t3.is_contiguous()):
            raise ValueError(f"{name}: expected contiguous [nfields, {self.ncol}, {pitch}], got {tuple(t.shape)}")
        return t3

    def setup(self, x1, x2):
        """LinInterp::setup for every column (impl :21-46, :160-186)."""
        self._coord(x1, self.pitch1, "x1")
        self._coord(x2, self.pitch2, "x2")
        fn = f"ekat_b200_lin_interp_setup_{_SFX[self.dtype]}"
        check(fn, getattr(lib(), fn)(x1.data_ptr(), x2.data_ptr(), self.indx_map.data_ptr(), self.ncol, self.km1,
                                     self.km2, self.pitch1, self.pitch2, _stream()))

    def lin_interp(self, x1, x2, y1, y2):
        """LinInterp::lin_interp for every column, using the map from setup (impl :48-158)."""
        self._coord(x1, self.pitch1, "x1")
        self._coord(x2, self.pitch2, "x2")
        y1f, y2f = self._field(y1, self.pitch1, "y1"), self._field(y2, self.pitch2, "y2")
        if y1f.shape[0] != y2f.shape[0]:
            raise ValueError("y1 / y2 field counts differ")
        fn = f"ekat_b200_lin_interp_apply_{_SFX[self.dtype]}"
        check(fn, getattr(lib(), fn)(x1.data_ptr(), x2.data_ptr(), y1f.data_ptr(), y2f.data_ptr(),
                                     self.indx_map.data_ptr(), self.ncol, self.km1, self.km2, self.pitch1, self.pitch2,
                                     y1f.shape[0], self.ncol * self.pitch1, self.ncol * self.pitch2, _stream()))
        return y2

    def run(self, x1, x2, y1, y2, write_map=False):
        """setup + lin_interp fused in one kernel; the index map is only written if write_map."""
        self._coord(x1, self.pitch1, "x1")
        self._coord(x2, self.pitch2, "x2")
        y1f, y2f = self._field(y1, self.pitch1, "y1"), self._field(y2, self.pitch2, "y2")
        if y1f.shape[0] != y2f.shape[0]:
            raise ValueError("y1 / y2 field counts differ")
        fn = f"ekat_b200_lin_interp_{_SFX[self.dtype]}"
        check(fn, getattr(lib(), fn)(x1.data_ptr(), x2.data_ptr(), y1f.data_ptr(), y2f.data_ptr(),
                                     self.indx_map.data_ptr() if write_map else None, self.ncol, self.km1, self.km2,
                                     self.pitch1, self.pitch2, y1f.shape[0], self.ncol * self.pitch1,
                                     self.ncol * self.pitch2, _stream()))
        return y2


def fill(dtype, ncol, km1, km2, pack_size=1, col0=0, seed=20260411, device="cuda", lev_major=False, out=None):
    """Device-side synthetic LinInterp inputs: returns x1, y1, x2.
    Reference layout [ncol, npack*N] (pack padding lanes zero), or level-major [km, ncol] if lev_major.
    out = (x1, y1, x2) from an earlier call of the same shape family: refill the first ncol columns in place."""
    if out is not None:
        x1, y1, x2 = out
        p1, p2 = (x1.shape[1], x2.shape[1])
    elif lev_major:
        x1 = torch.zeros((km1, ncol), dtype=dtype, device=device)
        y1 = torch.zeros((km1, ncol), dtype=dtype, device=device)
        x2 = torch.zeros((km2, ncol), dtype=dtype, device=device)
        p1 = p2 = ncol
    else:
        p1, p2 = npack(km1, pack_size) * pack_size, npack(km2, pack_size) * pack_size
        x1 = torch.zeros((ncol, p1), dtype=dtype, device=device)
        y1 = torch.zeros((ncol, p1), dtype=dtype, device=device)
        x2 = torch.zeros((ncol, p2), dtype=dtype, device=device)
    fn = f"ekat_b200_fill_lin_interp_{_SFX[dtype]}"
    check(fn, getattr(lib(), fn)(x1.data_ptr(), y1.data_ptr(), x2.data_ptr(), ncol, col0, km1, km2, p1, p2,
                                 1 if lev_major else 0, seed, _stream()))
    return x1, y1, x2


def tridiag_lin_interp(dl, d, du, B, x1, x2, Y2, X=None, ncol=None):
    """Fused shared-A multi-RHS Thomas solve + LinInterp of every solution (BASELINE config 4).
    Level-major operands with one column pitch: dl,d,du,x1 [nlev,pitch]; B [nrhs,nlev,pitch]; x2 [km2,pitch];
    Y2 [nrhs,km2,pitch]; X (optional) [nrhs,nlev,pitch]."""
    nrhs, nlev, pitch = B.shape
    km2 = x2.shape[0]
    ncol = pitch if ncol is None else ncol
    for t in (dl, d, du, x1):
        if t.shape != (nlev, pitch) or not t.is_contiguous():
            raise ValueError("dl,d,du,x1 must be contiguous [nlev,pitch]")
    if x2.shape != (km2, pitch) or Y2.shape != (nrhs, km2, pitch) or not (x2.is_contiguous() and Y2.is_contiguous()
                                                                           and B.is_contiguous()):
        raise ValueError("x2 [km2,pitch], Y2 [nrhs,km2,pitch], B [nrhs,nlev,pitch] must be contiguous")
    if X is not None and (X.shape != B.shape or not X.is_contiguous()):
        raise ValueError("X must match B")
    fn = f"ekat_b200_tridiag_lin_interp_{_SFX[B.dtype]}"
    check(fn, getattr(lib(), fn)(dl.data_ptr(), d.data_ptr(), du.data_ptr(), B.data_ptr(), x1.data_ptr(), x2.data_ptr(),
                                 Y2.data_ptr(), X.data_ptr() if X is not None else None, ncol, nlev, km2, nrhs, pitch,
                                 _stream()))
    return Y2


def lin_interp_lm(x1, x2, y1, y2, ncol=None):
    """ekat::LinInterp setup + lin_interp on level-major operands: x1 [km1,pitch], x2 [km2,pitch],
    y1 [nf,km1,pitch] (or [km1,pitch]), y2 [nf,km2,pitch] (or [km2,pitch]); contiguous, columns fastest."""
    y13 = y1 if y1.dim() == 3 else y1.unsqueeze(0)
    y23 = y2 if y2.dim() == 3 else y2.unsqueeze(0)
    nf, km1, pitch = y13.shape
    km2 = x2.shape[0]
    ncol = pitch if ncol is None else ncol
    if x1.shape != (km1, pitch) or x2.shape != (km2, pitch) or y23.shape != (nf, km2, pitch):
        raise ValueError("x1 [km1,pitch], x2 [km2,pitch], y1 [nf,km1,pitch], y2 [nf,km2,pitch]")
    for t in (x1, x2, y13, y23):
        if not (t.is_cuda and t.is_contiguous() and t.dtype == x1.dtype):
            raise ValueError("lin_interp_lm needs contiguous CUDA tensors of one dtype (there is no CPU path)")
    fn = f"ekat_b200_lin_interp_lm_{_SFX[x1.dtype]}"
    check(fn, getattr(lib(), fn)(x1.data_ptr(), x2.data_ptr(), y13.data_ptr(), y23.data_ptr(), ncol, km1, km2, pitch, nf,
                                 _stream()))
    return y2
