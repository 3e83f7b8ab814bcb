"""ctypes binding of libc2gpu.so -- the host-side mirror of the Fortran ``gpu_interface`` module.

Every call goes through the C ABI declared in include/c2gpu.h (the same symbols critic2's Fortran would bind with
``bind(c)``).  There is no CPU fallback: if the shared library is missing or no B200 is visible, calls raise.

Field objects mirror the reference's operator for this path, ``field%grd`` lifted to whole grids
(/root/reference/src/fieldmod@proc.f90:613-842): ``WfnField`` = type_wfn, ``PromolField`` = type_promol.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("C2GPU_LIB") or os.path.join(_HERE, "libc2gpu.so")   # C2GPU_LIB: A/B builds in development

PROPS = ("f", "gx", "gy", "gz", "gmod", "hxx", "hxy", "hxz", "hyy", "hyz", "hzz", "lap")  # rhoplot@proc.f90:488-523
RES_LEN = 16

dp = C.POINTER(C.c_double)
ip = C.POINTER(C.c_int)
_lib = None


class C2GpuError(RuntimeError):
    pass


class Unsupported(C2GpuError):
    """status -1: the caller should keep its own CPU loop."""


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise C2GpuError(f"{LIB_PATH} is missing: build it with `make -C critic2_b200/csrc` "
                             "(python __graft_entry__.py build); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        L.c2gpu_init.argtypes = [C.c_int, ip]
        L.c2gpu_init_devices.argtypes = [ip, C.c_int]
        L.c2gpu_last_error.argtypes = [C.c_char_p, C.c_int]
        L.c2gpu_launch_count.restype = C.c_long
        L.c2gpu_set_option.argtypes = [C.c_char_p, C.c_int]
        L.c2gpu_destroy.argtypes = [C.c_void_p]
        L.c2gpu_field_set_core.argtypes = [C.c_void_p, C.c_void_p]
        L.c2gpu_wfn_create.argtypes = [C.c_int, dp, C.c_int, ip, ip, dp, dp, C.c_int, C.c_int, dp, dp, C.c_int, C.c_int,
                                       C.c_int, ip, ip, dp, dp, dp, C.c_int, dp, dp, C.POINTER(C.c_void_p)]
        L.c2gpu_promol_create.argtypes = [C.c_int, C.c_int, dp, ip, ip, C.c_int, ip, ip, dp, dp, dp, dp, dp, dp, dp,
                                          C.c_int, dp, dp, dp, dp, C.c_char_p, C.POINTER(C.c_void_p)]
        L.c2gpu_eval_grid.argtypes = [C.c_void_p, dp, dp, ip, C.c_int, C.c_int, C.c_uint, C.POINTER(dp)]
        L.c2gpu_eval_grid_dev.argtypes = [C.c_void_p, dp, dp, ip, C.c_int, C.c_int, C.c_uint, C.POINTER(C.c_void_p),
                                          C.c_int, C.c_int, C.c_void_p]
        L.c2gpu_eval_grid_dev_part.argtypes = [C.c_void_p, dp, dp, ip, C.c_int, C.c_int, C.c_uint,
                                               C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_void_p]
        L.c2gpu_eval_grid_dev_part_rows.argtypes = L.c2gpu_eval_grid_dev_part.argtypes
        L.c2gpu_layer_height.argtypes = [C.c_void_p]
        L.c2gpu_gather_part_host.argtypes = [C.c_void_p, ip, C.c_int, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                             C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.c2gpu_gather_part_host_range.argtypes = [C.c_void_p, ip, C.c_int, C.c_int, C.POINTER(C.c_void_p),
                                                   C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                                   C.c_void_p]
        L.c2gpu_gather_wait.argtypes = [C.c_void_p]
        L.c2gpu_eval_grid_dev_part_range.argtypes = [C.c_void_p, dp, dp, ip, C.c_int, C.c_int, C.c_uint, C.POINTER(C.c_void_p),
                                                     C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.c2gpu_nci_grid_dev_part.argtypes = [C.c_void_p, C.c_void_p, dp, dp, ip, C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.c_int, C.c_int, C.c_void_p]
        L.c2gpu_nci_grid.argtypes = [C.c_void_p, C.c_void_p, dp, dp, ip, dp, dp, dp]
        L.c2gpu_nci_grid_dev.argtypes = [C.c_void_p, C.c_void_p, dp, dp, ip, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_int, C.c_int, C.c_void_p]
        L.c2gpu_nci_cutoffs.argtypes = [C.c_long, dp, dp, dp, C.c_double, C.c_double, C.c_double, dp, C.c_int,
                                        C.c_double, C.c_double, dp, C.POINTER(C.c_long), dp, C.c_long]
        L.c2gpu_nci_cutoffs_dev.argtypes = [C.c_long, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double,
                                            C.c_double, dp, C.c_int, C.c_double, dp, C.POINTER(C.c_long), C.c_void_p,
                                            C.c_long, C.c_void_p]
        L.c2gpu_nci_cutoffs_dev_noalloc.argtypes = L.c2gpu_nci_cutoffs_dev.argtypes
        L.c2gpu_nci_fragments.argtypes = [C.c_int, C.POINTER(C.c_void_p), dp, dp, ip, dp]
        L.c2gpu_nci_fragments_dev.argtypes = [C.c_int, C.POINTER(C.c_void_p), dp, dp, ip, C.c_void_p, C.c_int, C.c_int,
                                              C.c_void_p]
        L.c2gpu_nci_fragments_dev_part.argtypes = [C.c_int, C.POINTER(C.c_void_p), dp, dp, ip, C.c_void_p, C.c_int,
                                                   C.c_int, C.c_int, C.c_void_p]
        L.c2gpu_nci_cutoffs_frag.argtypes = [C.c_long, dp, dp, dp, C.c_int, dp, C.c_double, C.c_double, C.c_double,
                                             C.c_double, C.c_double, dp, C.c_int, C.c_double, C.c_double, dp,
                                             C.POINTER(C.c_long), dp, C.c_long]
        L.c2gpu_nci_cutoffs_frag_dev.argtypes = [C.c_long, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                                 C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, dp,
                                                 C.c_int, C.c_double, dp, C.POINTER(C.c_long), C.c_void_p, C.c_long,
                                                 C.c_void_p]
        L.c2gpu_cube_text_size.restype = C.c_long
        L.c2gpu_cube_text_size.argtypes = [ip, C.c_int]
        L.c2gpu_cube_text.argtypes = [dp, ip, C.c_int, C.c_int, C.c_char_p]
        L.c2gpu_cube_text_dev.argtypes = [C.c_void_p, ip, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.c2gpu_dat_text_size.restype = C.c_long
        L.c2gpu_dat_text_size.argtypes = [C.c_long]
        L.c2gpu_dat_text.argtypes = [C.c_long, dp, C.c_char_p]
        L.c2gpu_promol_grid.argtypes = [C.c_void_p, ip, dp]
        L.c2gpu_promol_grid_dev.argtypes = [C.c_void_p, ip, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.c2gpu_eval_points.argtypes = [C.c_void_p, C.c_long, dp, C.c_int, C.c_int, dp]
        L.c2gpu_eval_points_gkin.argtypes = [C.c_void_p, C.c_long, dp, C.c_int, dp]
        L.c2gpu_eval_points_mo.argtypes = [C.c_void_p, C.c_long, dp, C.c_int, dp]
        L.c2gpu_wfn_count_active.argtypes = [C.c_void_p, dp, dp, ip, C.c_int, C.c_int, C.c_int, dp]
        L.c2gpu_promol_count_active.argtypes = [C.c_void_p, dp, dp, ip, C.c_int, dp]
        L.c2gpu_wfn_config.argtypes = [C.c_void_p, ip]
        L.c2gpu_fp64_peak.argtypes = [dp, dp]
        _lib = L
    return _lib


def _check(rc: int):
    if rc == 0:
        return
    buf = C.create_string_buffer(512)
    lib().c2gpu_last_error(buf, 512)
    msg = buf.value.decode(errors="replace")
    if rc == -1:
        raise Unsupported(msg)
    raise C2GpuError(f"c2gpu status {rc}: {msg}")


def _d(a):
    return a.ctypes.data_as(dp)


def _i(a):
    return a.ctypes.data_as(ip)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _colmajor(m):
    return np.ascontiguousarray(np.asarray(m, float).T).reshape(-1)


def init(ndev: int = 0) -> int:
    n = C.c_int(0)
    _check(lib().c2gpu_init(ndev, C.byref(n)))
    return n.value


def init_devices(devs) -> None:
    a = _i32(list(devs))
    _check(lib().c2gpu_init_devices(_i(a), len(a)))


def set_option(name: str, value: int) -> None:
    _check(lib().c2gpu_set_option(name.encode(), int(value)))


def fp64_peak():
    """(DMMA TFLOP/s, DFMA TFLOP/s) of the first device, measured now."""
    a, b = C.c_double(0), C.c_double(0)
    _check(lib().c2gpu_fp64_peak(C.byref(a), C.byref(b)))
    return a.value, b.value


def launch_count() -> int:
    return int(lib().c2gpu_launch_count())


def _grid_args(x0, xd, nn):
    return _f64(x0), _colmajor(xd), _i32(nn)


class _Field:
    """Common grid drivers of a field handle."""

    handle = None

    def close(self):
        if self.handle is not None and _lib is not None:
            _lib.c2gpu_destroy(self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- f%usecore / f%zpsp, fieldmod@proc.f90:154-157,184-199; used by grd :827-832 and grd0 :922-925 ------------
    def set_core(self, core: "PromolField | None"):
        """Attach (or detach with None) the core-density field built by ``core_field``: every driver of this field
        then adds the promolecular core density and its derivatives before the tail of grd."""
        _check(lib().c2gpu_field_set_core(self.handle, core.handle if core is not None else None))
        self._core = core   # keep it alive: the library does not own it

    # -- rhoplot_cube evaluation loop, rhoplot@proc.f90:582-631 -------------------------------------------------
    def cube(self, x0, xd, nn, props=PROPS, nder: int = 2, periodic: bool = False):
        """HOST arrays through the C ABI (copies inside). xd[:, i] is the i-th step vector.
        Returns {prop: array (nn3, nn2, nn1)} = lf(ix,iy,iz) with x fastest."""
        x0a, xda, nna = _grid_args(x0, xd, nn)
        names = [p for p in PROPS if p in props]
        mask = sum(1 << PROPS.index(p) for p in names)
        npt = int(nna[0]) * int(nna[1]) * int(nna[2])
        outs = [np.empty(npt) for _ in names]
        arr = (dp * len(outs))(*[_d(o) for o in outs])
        _check(lib().c2gpu_eval_grid(self.handle, _d(x0a), _d(xda), _i(nna), nder, int(periodic), mask, arr))
        return {p: o.reshape(nna[2], nna[1], nna[0]) for p, o in zip(names, outs)}

    def cube_dev(self, x0, xd, nn, outs, props, nder: int = 2, periodic: bool = False, iz0: int = 0, iz1=None,
                 stream: int = 0):
        """DEVICE arrays (torch CUDA float64 tensors of nn1*nn2*nn3 elements, one per property in PROPS order)."""
        x0a, xda, nna = _grid_args(x0, xd, nn)
        names = [p for p in PROPS if p in props]
        mask = sum(1 << PROPS.index(p) for p in names)
        assert len(outs) == len(names)
        ptrs = (C.c_void_p * len(outs))(*[C.c_void_p(t.data_ptr()) for t in outs])
        _check(lib().c2gpu_eval_grid_dev(self.handle, _d(x0a), _d(xda), _i(nna), nder, int(periodic), mask, ptrs,
                                         iz0, int(nna[2]) if iz1 is None else iz1, C.c_void_p(stream)))

    def cube_dev_part(self, x0, xd, nn, outs, props, part: int, nparts: int, nder: int = 2, periodic: bool = False,
                      stream: int = 0, rows: bool = False):
        """Like cube_dev for the interleaved z layers (layer % nparts == part): the multi-GPU decomposition.
        rows=True interleaves rows of super-tiles instead of whole layers (finer balance; gather by reduction)."""
        x0a, xda, nna = _grid_args(x0, xd, nn)
        names = [p for p in PROPS if p in props]
        mask = sum(1 << PROPS.index(p) for p in names)
        ptrs = (C.c_void_p * len(outs))(*[C.c_void_p(t.data_ptr()) for t in outs])
        fn = lib().c2gpu_eval_grid_dev_part_rows if rows else lib().c2gpu_eval_grid_dev_part
        _check(fn(self.handle, _d(x0a), _d(xda), _i(nna), nder, int(periodic), mask, ptrs, part, nparts, C.c_void_p(stream)))

    def cube_dev_part_range(self, x0, xd, nn, outs, props, part: int, nparts: int, iz0: int, iz1: int, nder: int = 2,
                            periodic: bool = False, stream: int = 0, rows: bool = True):
        """cube_dev_part on the slab [iz0, iz1) of z planes (units counted from iz0)."""
        x0a, xda, nna = _grid_args(x0, xd, nn)
        names = [p for p in PROPS if p in props]
        mask = sum(1 << PROPS.index(p) for p in names)
        ptrs = (C.c_void_p * len(outs))(*[C.c_void_p(t.data_ptr()) for t in outs])
        _check(lib().c2gpu_eval_grid_dev_part_range(self.handle, _d(x0a), _d(xda), _i(nna), nder, int(periodic), mask, ptrs,
                                                    part, nparts, int(rows), iz0, iz1, C.c_void_p(stream)))

    def gather_part_host_range(self, nn, dev_arrays, host_ptrs, part: int, nparts: int, c0: int, c1: int, rows: bool = True,
                               nci_order: bool = False, async_: bool = True, stream: int = 0):
        na = _i32(nn)
        d = (C.c_void_p * len(dev_arrays))(*[C.c_void_p(t.data_ptr()) for t in dev_arrays])
        hp = (C.c_void_p * len(host_ptrs))(*[C.c_void_p(int(p)) for p in host_ptrs])
        _check(lib().c2gpu_gather_part_host_range(self.handle, _i(na), int(nci_order), len(dev_arrays), d, hp, part, nparts,
                                                  int(rows), c0, c1, int(async_), C.c_void_p(stream)))

    def gather_wait(self):
        _check(lib().c2gpu_gather_wait(self.handle))

    def layer_height(self) -> int:
        return int(lib().c2gpu_layer_height(self.handle))

    def gather_part_host(self, nn, dev_arrays, host_ptrs, part: int, nparts: int, rows: bool = False,
                         nci_order: bool = False, stream: int = 0):
        """Copy the blocks `part` owns from full-size DEVICE tensors into full-size HOST arrays (addresses in host_ptrs;
        one shared page-locked buffer when every rank passes the same memory): the gather to the writer rank."""
        na = _i32(nn)
        d = (C.c_void_p * len(dev_arrays))(*[C.c_void_p(t.data_ptr()) for t in dev_arrays])
        hp = (C.c_void_p * len(host_ptrs))(*[C.c_void_p(int(p)) for p in host_ptrs])
        _check(lib().c2gpu_gather_part_host(self.handle, _i(na), int(nci_order), len(dev_arrays), d, hp, part, nparts,
                                            int(rows), C.c_void_p(stream)))

    # -- nciplot evaluation loop, nci@proc.f90:499-561 ----------------------------------------------------------
    def nci(self, x0, xmat, nstep, promol: "_Field | None" = None):
        """Returns crho, cgrad[, rhoat] as (n1, n2, n3) arrays, z fastest (the reference's crho(k,j,i))."""
        x0a, xda, nna = _grid_args(x0, xmat, nstep)
        npt = int(nna[0]) * int(nna[1]) * int(nna[2])
        crho, cgrad = np.empty(npt), np.empty(npt)
        rhoat = np.empty(npt) if promol is not None else None
        _check(lib().c2gpu_nci_grid(self.handle, promol.handle if promol is not None else None, _d(x0a), _d(xda),
                                    _i(nna), _d(crho), _d(cgrad), _d(rhoat) if rhoat is not None else None))
        shp = (nna[0], nna[1], nna[2])
        return crho.reshape(shp), cgrad.reshape(shp), (rhoat.reshape(shp) if rhoat is not None else None)

    def nci_dev(self, x0, xmat, nstep, crho, cgrad, rhoat=None, promol: "_Field | None" = None, i0: int = 0, i1=None,
                stream: int = 0):
        x0a, xda, nna = _grid_args(x0, xmat, nstep)
        _check(lib().c2gpu_nci_grid_dev(self.handle, promol.handle if promol is not None else None, _d(x0a), _d(xda),
                                        _i(nna), C.c_void_p(crho.data_ptr()), C.c_void_p(cgrad.data_ptr()),
                                        C.c_void_p(rhoat.data_ptr()) if rhoat is not None else None, i0,
                                        int(nna[0]) if i1 is None else i1, C.c_void_p(stream)))

    def nci_dev_part(self, x0, xmat, nstep, crho, cgrad, part: int, nparts: int, rhoat=None,
                     promol: "_Field | None" = None, stream: int = 0):
        x0a, xda, nna = _grid_args(x0, xmat, nstep)
        _check(lib().c2gpu_nci_grid_dev_part(self.handle, promol.handle if promol is not None else None, _d(x0a),
                                             _d(xda), _i(nna), C.c_void_p(crho.data_ptr()),
                                             C.c_void_p(cgrad.data_ptr()),
                                             C.c_void_p(rhoat.data_ptr()) if rhoat is not None else None, part, nparts,
                                             C.c_void_p(stream)))

    # -- field%grd on a point list, fieldmod@proc.f90:613-842 ---------------------------------------------------
    def points(self, pts, nder: int = 2, periodic: bool = True):
        pts = _f64(pts).reshape(-1, 3)
        res = np.empty((pts.shape[0], RES_LEN))
        _check(lib().c2gpu_eval_points(self.handle, pts.shape[0], _d(pts), nder, int(periodic), _d(res)))
        return res


class WfnField(_Field):
    """type_wfn field: the flat molwfn arrays (critic2_b200.wfn.Molwfn) placed in a molecular cell
    (critic2_b200.structure.Structure); wfn_private.f90:83-131."""

    def __init__(self, f, st):
        if f.dran is None:
            from .wfn import complete_struct
            complete_struct(f)
        self.f, self.st = f, st
        h = C.c_void_p()
        ne = f.nedf
        nmoall = f.cmo.shape[1]
        _check(lib().c2gpu_wfn_create(
            f.natom, _d(_f64(st.r)), f.npri, _i(_i32(f.icenter)), _i(_i32(f.itype)), _d(_f64(f.e)), _d(_f64(f.dran)),
            f.nmoocc, nmoall, _d(_f64(f.cmo)), _d(_f64(f.occ)), f.wfntyp, f.nalpha, ne,
            _i(_i32(f.icenter_edf)) if ne else None, _i(_i32(f.itype_edf)) if ne else None,
            _d(_f64(f.e_edf)) if ne else None, _d(_f64(f.c_edf)) if ne else None,
            _d(_f64(f.dran_edf)) if ne else None, int(st.ismolecule), _d(_colmajor(st.m_x2c)), _d(_colmajor(st.m_c2x)),
            C.byref(h)))
        self.handle = h

    # -- res%gkin of grd (rho2, wfn_private@proc.F90:2104-2106,2209): mesh%fill's im_gkin ----------------------------
    def gkin_points(self, pts, periodic: bool = True):
        pts = _f64(pts).reshape(-1, 3)
        out = np.empty(pts.shape[0])
        _check(lib().c2gpu_eval_points_gkin(self.handle, pts.shape[0], _d(pts), int(periodic), _d(out)))
        return out

    # -- grd(nder=-1, fder) -> calculate_mo (wfn_private@proc.F90:2447-2553): mesh%fill's prop > 100 --------------------
    def mo_field(self, imo: int) -> "MoField":
        """One-MO companion handle for MO number imo (1-based, <= nmoall); cached per MO."""
        cache = self.__dict__.setdefault("_mo", {})
        if imo not in cache:
            cache[imo] = MoField(self.f, self.st, imo)
        return cache[imo]

    def mo_points(self, pts, imo: int, periodic: bool = True):
        return self.mo_field(imo).values(pts, periodic)

    # -- the spin and valence channels of rho2 (wfn_private@proc.F90:2091-2147, 2159-2200; grd :763-786) ---------------------
    def channel(self, spin: int = 0, valence: bool = False) -> "WfnField":
        """Companion field whose density is one channel of rho2: spin = 1 / 2 -> rho(2) / rho(3) (res%fspin, gfmodspin,
        lapspin, gkinspin: the alpha / beta MOs of a UHF wavefunction plus HALF the EDF core, :2184-2198; half of everything
        otherwise, :2159-2182); valence = True -> the channel without the EDF core (res%fval, gfmodval, del2fval and the
        *spinval members).  Same kernels, same primitives: only the occupations and the EDF coefficients of the handle
        differ, so every driver (CUBE, NCIPLOT, point lists, gkin) works on a channel.  Cached per (spin, valence)."""
        import copy
        from .wfn import WFN_UHF
        if spin not in (0, 1, 2):
            raise C2GpuError("channel: spin must be 0 (total), 1 (alpha) or 2 (beta)")
        if spin == 0 and not valence:
            return self
        cache = self.__dict__.setdefault("_chan", {})
        key = (spin, bool(valence))
        if key not in cache:
            f = self.f
            g = copy.copy(f)
            occ = np.array(f.occ, float)
            if spin:
                if f.wfntyp == WFN_UHF:
                    alpha = np.arange(f.nmoocc) < f.nalpha           # imo <= nalpha, :2096
                    occ = np.where(alpha if spin == 1 else ~alpha, occ, 0.0)
                else:
                    occ = 0.5 * occ                                   # rho(2) = rho(3) = rho(1) / 2, :2163-2164
            g.occ = occ
            if valence or f.nedf == 0:
                g.icenter_edf, g.itype_edf = np.zeros(0, np.int32), np.zeros(0, np.int32)
                g.e_edf, g.c_edf, g.dran_edf = np.zeros(0), np.zeros(0), np.zeros(0)
            elif spin:
                g.c_edf = 0.5 * np.asarray(f.c_edf, float)           # rho(2:3) += rhoc / 2, :2186
            cache[key] = WfnField(g, self.st)
        return cache[key]

    def config(self):
        a = np.zeros(8, np.int32)
        _check(lib().c2gpu_wfn_config(self.handle, _i(a)))
        return dict(cfg=int(a[0]), points_per_tile=int(a[1]), mo_block=int(a[2]), nblk=int(a[3]), threads=int(a[4]),
                    kchunk=int(a[5]), nshell=int(a[6]), grid=int(a[7]))

    def count_active(self, x0, xd, nn, periodic: bool = False, iz0: int = 0, iz1=None) -> float:
        x0a, xda, nna = _grid_args(x0, xd, nn)
        s = C.c_double(0)
        _check(lib().c2gpu_wfn_count_active(self.handle, _d(x0a), _d(xda), _i(nna), int(periodic), iz0,
                                            int(nna[2]) if iz1 is None else iz1, C.byref(s)))
        return s.value


class MoField(_Field):
    """The one-MO wavefunction handle c2gpu_eval_points_mo evaluates: same primitives, cmo = row imo (1-based)."""

    def __init__(self, f, st, imo: int):
        if not 1 <= imo <= f.cmo.shape[1]:
            raise C2GpuError("Invalid molecular orbital")        # calculate_mo :2531-2532
        h = C.c_void_p()
        row = np.ascontiguousarray(f.cmo[:, imo - 1])
        one = np.ones(1)
        _check(lib().c2gpu_wfn_create(
            f.natom, _d(_f64(st.r)), f.npri, _i(_i32(f.icenter)), _i(_i32(f.itype)), _d(_f64(f.e)), _d(_f64(f.dran)),
            1, 1, _d(row), _d(one), 0, 1, 0, None, None, None, None, None, int(st.ismolecule), _d(_colmajor(st.m_x2c)),
            _d(_colmajor(st.m_c2x)), C.byref(h)))
        self.handle = h
        self.imo = imo

    def values(self, pts, periodic: bool = True):
        pts = _f64(pts).reshape(-1, 3)
        out = np.empty(pts.shape[0])
        _check(lib().c2gpu_eval_points_mo(self.handle, pts.shape[0], _d(pts), int(periodic), _d(out)))
        return out


class PromolField(_Field):
    """type_promol field: environment atoms + radial tables. tables: {Z: dict(a,b,r,f,fp,fpp)}."""

    def __init__(self, st, tables, fragmask=None):
        from .structure import CUTRAD
        self.st = st
        zspc, iss = st.species()
        r_env, cidx = st.environment_atoms()
        ispc = iss[cidx - 1]
        ngrid, aa, bb, cut = [], [], [], []
        tr, tf, tfp, tfpp = [], [], [], []
        for z in zspc:
            t = tables.get(int(z))
            cut.append(float(CUTRAD[int(z) - 1]))
            if t is None:
                ngrid.append(0); aa.append(0.0); bb.append(1.0)
                continue
            ngrid.append(len(t["r"])); aa.append(float(t["a"])); bb.append(float(t["b"]))
            tr.append(_f64(t["r"])); tf.append(_f64(t["f"])); tfp.append(_f64(t["fp"])); tfpp.append(_f64(t["fpp"]))
        cat = lambda x: np.concatenate(x) if x else np.zeros(1)
        h = C.c_void_p()
        frag = None if fragmask is None else np.ascontiguousarray(fragmask, np.uint8).tobytes()
        _check(lib().c2gpu_promol_create(
            r_env.shape[0], st.ncell, _d(_f64(r_env)), _i(_i32(ispc)), _i(_i32(cidx)), len(zspc), _i(_i32(zspc)),
            _i(_i32(ngrid)), _d(_f64(aa)), _d(_f64(bb)), _d(_f64(cut)), _d(cat(tr)), _d(cat(tf)), _d(cat(tfp)),
            _d(cat(tfpp)), int(st.ismolecule), _d(_colmajor(st.m_x2c)), _d(_colmajor(st.m_c2x)),
            _d(_colmajor(st.m_xr2c)), _d(_colmajor(st.m_c2xr)), frag, C.byref(h)))
        self.handle = h
        self.nenv = int(r_env.shape[0])

    # -- crystal%promolecular_grid, crystalmod@proc.f90:5118-5158 ------------------------------------------------
    def promolecular_grid(self, n):
        na = _i32(n)
        out = np.empty(int(na[0]) * int(na[1]) * int(na[2]))
        _check(lib().c2gpu_promol_grid(self.handle, _i(na), _d(out)))
        return out.reshape(na[2], na[1], na[0])

    def promolecular_grid_dev(self, n, out, iz0: int = 0, iz1=None, stream: int = 0):
        na = _i32(n)
        _check(lib().c2gpu_promol_grid_dev(self.handle, _i(na), C.c_void_p(out.data_ptr()), iz0,
                                           int(na[2]) if iz1 is None else iz1, C.c_void_p(stream)))

    def count_active(self, x0, xd, nn, periodic: bool = False) -> float:
        x0a, xda, nna = _grid_args(x0, xd, nn)
        s = C.c_double(0)
        _check(lib().c2gpu_promol_count_active(self.handle, _d(x0a), _d(xda), _i(nna), int(periodic), C.byref(s)))
        return s.value


def core_field(st, core_tables, zpsp) -> PromolField:
    """The promolecular field of the CORE densities cgrid(iz, zpsp(iz)) (grid1mod@proc.f90:139-164) over the structure's
    environment.  core_tables: {(Z, q): dict(a,b,r,f,fp,fpp)} as read_db(Z, q) builds them; zpsp: {Z: q}.  Elements with
    zpsp <= 0 or without core electrons (Z - q <= 0, grid1mod@proc.f90:100) get no table and contribute nothing
    (environmod@proc.f90:1299)."""
    tabs = {}
    for z, q in zpsp.items():
        if q > 0 and z - q > 0:
            tabs[int(z)] = core_tables[(int(z), int(q))]
    return PromolField(st, tabs)


class FragmentField(_Field):
    """One NCIPLOT fragment (nci@proc.f90:534-550): a promolecular handle over the fragment's own atoms only -- no
    periodic images, cutoff g%rmax (cutrad = 1e30), molecule frame with identity matrices so the point is never folded.
    xyz: (nat, 3) Cartesian bohr (the reference's fr%at(:)%r); z: atomic numbers; tables as for PromolField."""

    def __init__(self, xyz, z, tables):
        xyz = _f64(xyz).reshape(-1, 3)
        z = _i32(z).reshape(-1)
        nat = xyz.shape[0]
        zspc = np.unique(z)
        ispc = _i32(np.searchsorted(zspc, z) + 1)
        cidx = _i32(np.arange(1, nat + 1))
        ngrid, aa, bb, tr, tf = [], [], [], [], []
        for zz in zspc:
            t = tables.get(int(zz))
            if t is None:                      # agrid(iz) not initialised: the atom is skipped (:541)
                ngrid.append(0); aa.append(0.0); bb.append(1.0)
                continue
            ngrid.append(len(t["r"])); aa.append(float(t["a"])); bb.append(float(t["b"]))
            tr.append(_f64(t["r"])); tf.append(_f64(t["f"]))
        cat = lambda x: np.concatenate(x) if x else np.zeros(1)
        r_, f_ = cat(tr), cat(tf)
        eye = np.eye(3)
        h = C.c_void_p()
        cut = np.full(len(zspc), 1e30)
        _check(lib().c2gpu_promol_create(nat, nat, _d(xyz), _i(ispc), _i(cidx), len(zspc), _i(_i32(zspc)), _i(_i32(ngrid)),
                                         _d(_f64(aa)), _d(_f64(bb)), _d(cut), _d(r_), _d(f_), _d(f_), _d(f_), 1,
                                         _d(_colmajor(eye)), _d(_colmajor(eye)), _d(_colmajor(eye)), _d(_colmajor(eye)),
                                         None, C.byref(h)))
        self.handle = h
        self.nat = nat


def nci_fragments(frags, x0, xmat, nstep):
    """rhofrag(k,j,i,ifr) on HOST arrays: returns (nfrag, n1, n2, n3), z fastest."""
    x0a, xda, nna = _grid_args(x0, xmat, nstep)
    npt = int(nna[0]) * int(nna[1]) * int(nna[2])
    out = np.empty(len(frags) * npt)
    hs = (C.c_void_p * len(frags))(*[f.handle for f in frags])
    _check(lib().c2gpu_nci_fragments(len(frags), hs, _d(x0a), _d(xda), _i(nna), _d(out)))
    return out.reshape(len(frags), nna[0], nna[1], nna[2])


def nci_fragments_dev(frags, x0, xmat, nstep, rhofrag, part: int = 0, nparts: int = 1, layer_height: int = 8,
                      stream: int = 0):
    """rhofrag on a DEVICE tensor of nfrag*n1*n2*n3 doubles; interleaved x-layers like nci_dev_part."""
    x0a, xda, nna = _grid_args(x0, xmat, nstep)
    hs = (C.c_void_p * len(frags))(*[f.handle for f in frags])
    _check(lib().c2gpu_nci_fragments_dev_part(len(frags), hs, _d(x0a), _d(xda), _i(nna), C.c_void_p(rhofrag.data_ptr()),
                                              part, nparts, layer_height, C.c_void_p(stream)))


def nci_cutoffs(crho, cgrad, rhoat=None, rhocut=0.2, dimcut=2.0, rhoplot=0.1, srhorange=(-1e30, 1e30), onlyneg=False,
                rho_void=-1.0, omega=1.0, want_dat=True, rhofrag=None, rhoparam=0.95, rhoparam2=0.75):
    """nciplot cutoff pass (nci@proc.f90:569-604) on HOST arrays; cgrad is modified in place.
    rhofrag: None or (nfrag, npts) fragment densities (then rhoat is required, :579-582).
    Returns (sums[3], dat (ndat, 2), ndat)."""
    npts = crho.size
    sums = np.zeros(3)
    nd = C.c_long(0)
    dat = np.zeros((npts if want_dat else 0, 2))
    sr = _f64(srhorange)
    assert crho.flags.c_contiguous and cgrad.flags.c_contiguous
    nfrag = 0 if rhofrag is None else int(np.asarray(rhofrag).shape[0])
    rf = None if rhofrag is None else _f64(rhofrag).reshape(nfrag, -1)
    _check(lib().c2gpu_nci_cutoffs_frag(npts, _d(crho), _d(cgrad), _d(rhoat) if rhoat is not None else None, nfrag,
                                        _d(rf) if rf is not None else None, rhoparam, rhoparam2, rhocut, dimcut, rhoplot,
                                        _d(sr), int(onlyneg), rho_void, omega, _d(sums), C.byref(nd),
                                        _d(dat) if want_dat else None, npts if want_dat else 0))
    return sums, dat[:nd.value].copy(), nd.value


# -- text bodies of the output files: writegrid_cube (crystalmod@proc.f90:5026-5036), write_cube_body (nci@proc.f90:867-883)
FMT_PRECISE, FMT_CUBE, FMT_NCI = 0, 1, 2


def cube_text(g, fmt: int = FMT_PRECISE, order: int = 0) -> bytes:
    """Formatted body of a cube file from a HOST array.  order 0: g indexed [iz][iy][ix] (lf(ix,iy,iz), x fastest),
    records over (ix, iy) as writegrid_cube; order 1: g indexed [i][j][k] (crho(k,j,i), z fastest), records over (i, j)
    as write_cube_body.  fmt: FMT_PRECISE '(6(1x,e22.14))', FMT_CUBE '(1p,6(1x,e12.5))', FMT_NCI '(6(1x,1p,e13.5e3))'."""
    g = _f64(g)
    assert g.ndim == 3
    n = _i32(g.shape[::-1] if order == 0 else g.shape)
    nb = lib().c2gpu_cube_text_size(_i(n), fmt)
    buf = C.create_string_buffer(nb)
    _check(lib().c2gpu_cube_text(_d(g), _i(n), fmt, order, buf))
    return buf.raw


def cube_text_dev(g, n, text, fmt: int = FMT_PRECISE, order: int = 0, stream: int = 0):
    """DEVICE version: g a CUDA float64 tensor, text a CUDA uint8 tensor of cube_text_size(n, fmt) bytes."""
    na = _i32(n)
    _check(lib().c2gpu_cube_text_dev(C.c_void_p(g.data_ptr()), _i(na), fmt, order, C.c_void_p(text.data_ptr()),
                                     C.c_void_p(stream)))


def cube_text_size(n, fmt: int = FMT_PRECISE) -> int:
    return int(lib().c2gpu_cube_text_size(_i(_i32(n)), fmt))


def dat_text(pairs) -> bytes:
    """'(1p,E15.7,E15.7)' lines of the NCIPLOT .dat file (nci@proc.f90:593-595) from (ndat, 2) pairs."""
    p = _f64(pairs).reshape(-1, 2)
    if p.shape[0] == 0:
        return b""
    buf = C.create_string_buffer(int(lib().c2gpu_dat_text_size(p.shape[0])))
    _check(lib().c2gpu_dat_text(p.shape[0], _d(p), buf))
    return buf.raw
