"""``FluxPlan`` -- Python handle on a ``cmb200_plan``: the grids, bands and absorption tables
of one ``RealGas`` configuration resident in HBM, plus the batched flux evaluation over any
number of independent columns (each column is one ``ny=1`` ``RealGas``, SURVEY.md finding 3).

Two call styles:

* ``flux(...)``        -- NumPy in, NumPy out (``cmb200_flux_host``; copies inside the call);
* ``flux_device(...)`` -- ``torch`` CUDA tensors in and out on the current stream
                          (``cmb200_flux``); torch is used for device memory only.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import FluxIO, as_f64, check, dptr

_PRECISION = {'f64': _lib.F64, 'fp64': _lib.F64, 'float64': _lib.F64,
              'f32': _lib.F32, 'fp32': _lib.F32, 'float32': _lib.F32}


def bands_from_nu_bands(nu, nu_bands):
    """Flatten the reference's ``nu_bands`` dict (real_gas.py:384-388: 'range' list of arrays,
    'centre', 'delta', 'sw') into bin runs (start, len) on ``nu``.  Each band must be a
    contiguous run of ``nu`` -- which is how get_wavenumber_bands builds them."""
    nu = np.asarray(nu, dtype=float)
    start, length = [], []
    for r in nu_bands['range']:
        r = np.asarray(r, dtype=float)
        idx = np.where(nu == r[0])[0]
        if idx.size == 0 or idx[0] + r.size > nu.size or not np.array_equal(nu[idx[0]:idx[0] + r.size], r):
            raise ValueError('band range is not a contiguous run of the wavenumber grid')
        start.append(int(idx[0]))
        length.append(int(r.size))
    return {'start': np.array(start, dtype=np.int32), 'len': np.array(length, dtype=np.int32),
            'centre': as_f64(nu_bands['centre']), 'delta': as_f64(nu_bands['delta']),
            'sw': np.asarray(nu_bands['sw']).astype(np.uint8)}


def segment_table(band_len, band_sw, bins_per_segment=32):
    """The kernel's unit of work sharing: runs of ``bins_per_segment`` bins inside a band, numbered
    band after band.  Returns (band index, is-long-wave flag) of every segment."""
    per_band = -(-np.asarray(band_len, dtype=np.int64) // bins_per_segment)
    seg_band = np.repeat(np.arange(len(per_band)), per_band)
    return seg_band, ~np.asarray(band_sw).astype(bool)[seg_band]


def shard_edges(is_lw, world_size, lw_cost=2.55):
    """Cut the segment list into ``world_size`` contiguous ranges of equal estimated cost (a
    long-wave segment also runs the emission recurrences and, being alone in its band, pays a
    replay-only pass).  Returns world_size + 1 edges; rank r owns [edges[r], edges[r + 1])."""
    cost = np.cumsum(np.where(is_lw, lw_cost, 1.0))
    inner = [int(np.searchsorted(cost, cost[-1] * r / world_size)) for r in range(1, world_size)]
    return [0] + inner + [int(np.size(is_lw))]


def wavenumber_shards(n_segments, world_size, min_segments=128):
    """How many ranks should share one set of columns: the largest divisor of ``world_size`` that
    leaves every shard at least ``min_segments`` segments (16 per warp of the 8-warp CTA -- below
    that the whole-segment split over the warps no longer balances); the remaining factor of
    world_size splits the columns, which costs no exchange at all."""
    best = 1
    for s in range(1, world_size + 1):
        if world_size % s == 0 and n_segments >= s * min_segments:
            best = s
    return best


class FluxPlan:
    def __init__(self, P, nu, bands, tables=None, precision='f64', device=0,
                 T_star=None, R_star=None, star_planet_dist=None):
        from . import constants as k
        lib = _lib.load()
        self._lib = lib
        self._h = C.c_void_p()
        self.P = as_f64(P).reshape(-1)
        self.nu = as_f64(nu).reshape(-1)
        self.L = self.P.size
        self.N = self.nu.size
        self.band_start = np.ascontiguousarray(bands['start'], dtype=np.int32)
        self.band_len = np.ascontiguousarray(bands['len'], dtype=np.int32)
        self.band_centre = as_f64(bands['centre'])
        self.band_delta = as_f64(bands['delta'])
        self.band_sw = np.ascontiguousarray(np.asarray(bands['sw']).astype(np.uint8))
        self.B = self.band_centre.size
        if not (self.band_start.size == self.band_len.size == self.band_delta.size == self.band_sw.size == self.B):
            raise ValueError('band arrays have different lengths')
        self.precision = precision
        self.device = int(device)
        self.M = 0
        cst = _lib.constants_struct()
        check(lib.cmb200_plan_create(
            C.byref(self._h), self.device, _PRECISION[precision], C.byref(cst), self.L, dptr(self.P),
            self.N, dptr(self.nu), self.B,
            self.band_start.ctypes.data_as(_lib.c_int32_p), self.band_len.ctypes.data_as(_lib.c_int32_p),
            dptr(self.band_centre), dptr(self.band_delta), self.band_sw.ctypes.data_as(_lib.c_uint8_p)))
        self.set_star(k.T_sun if T_star is None else T_star, k.R_sun if R_star is None else R_star,
                      k.AU if star_planet_dist is None else star_planet_dist)
        if tables is not None:
            self.set_tables(tables)

    # -- lifetime ---------------------------------------------------------------------------
    def close(self):
        if getattr(self, '_h', None) is not None and self._h.value:
            self._lib.cmb200_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- configuration ----------------------------------------------------------------------
    def set_star(self, T_star, R_star, dist):
        check(self._lib.cmb200_plan_set_star(self._h, float(T_star), float(R_star), float(dist)))
        self.T_star, self.R_star, self.dist = float(T_star), float(R_star), float(dist)

    def set_tables(self, tables):
        """tables: list of reference-format dicts {'p','T','nu','absorption_coef'}."""
        keep = []
        arr = (_lib.Table * len(tables))(*[_lib.table_struct(t, keep) for t in tables])
        check(self._lib.cmb200_plan_set_tables(self._h, len(tables), arr))
        self.M = len(tables)

    def set_band_range(self, lo, hi):
        check(self._lib.cmb200_plan_set_band_range(self._h, int(lo), int(hi)))
        self.band_lo, self.band_hi = int(lo), int(hi)

    def segments(self):
        """(bins_per_segment, band index of every segment, is-long-wave flag of every segment)."""
        bps, n = C.c_int32(), C.c_int64()
        check(self._lib.cmb200_plan_segments(self._h, C.byref(bps), C.byref(n)))
        seg_band, is_lw = segment_table(self.band_len, self.band_sw, bps.value)
        assert seg_band.size == n.value
        return bps.value, seg_band, is_lw

    def set_segment_range(self, lo, hi):
        check(self._lib.cmb200_plan_set_segment_range(self._h, int(lo), int(hi)))

    def shard_by_wavenumber(self, rank, world_size, lw_cost=2.55):
        """Restrict this plan to rank's share of the spectrum: contiguous segment ranges of
        equal estimated cost (a long-wave segment also runs the emission recurrences).  The
        outputs are then partial sums; add them across ranks."""
        _, _, is_lw = self.segments()
        edges = shard_edges(is_lw, world_size, lw_cost)
        self.set_segment_range(edges[rank], edges[rank + 1])
        return edges[rank], edges[rank + 1]

    @property
    def n_chains(self):
        n = C.c_int64()
        check(self._lib.cmb200_plan_info(self._h, None, None, None, None, None, None, C.byref(n)))
        return n.value

    # -- host (NumPy) path ------------------------------------------------------------------
    def flux(self, T, T_g, tau=None, tau_id=None, T_tau=None, q=None, albedo=0.3,
             outputs=('up', 'down', 'net', 'heating'), out=None):
        """Fluxes for n_col columns.  T [n_col, L-1] (or [L-1]), T_g [n_col] (or scalar).

        Optical-depth source: ``tau`` [L, N] or [n_tau, L, N] (+ ``tau_id`` [n_col]) -- or, when
        ``tau`` is None, the plan's tables with lookup temperatures ``T_tau`` [n_col, L] (None:
        isothermal at T_g) and specific humidities ``q`` [M, L] or [M, n_col, L].
        Returns a dict of NumPy arrays: up/down [n_col, L, B], net [n_col, L],
        heating [n_col, L-1], T_interface [n_col, L] -- as requested in ``outputs``; arrays given in
        ``out`` (e.g. page-locked ones, so that the copies overlap the kernels) are filled in place."""
        L, B = self.L, self.B
        T = as_f64(T)
        if T.ndim == 1:
            T = T.reshape(1, -1)
        n_col = T.shape[0]
        if T.shape != (n_col, L - 1):
            raise ValueError('T must have shape [n_col, %d], got %s' % (L - 1, T.shape))
        T_g = np.ascontiguousarray(np.broadcast_to(as_f64(T_g).reshape(-1), (n_col,)))
        io = FluxIO()
        io.n_col = n_col
        keep = [T, T_g]
        io.T, io.T_g = T.ctypes.data, T_g.ctypes.data
        if np.ndim(albedo) == 0:
            io.albedo, io.albedo_scalar = None, float(albedo)
        else:
            a = np.ascontiguousarray(np.broadcast_to(as_f64(albedo).reshape(-1), (n_col,)))
            keep.append(a)
            io.albedo = a.ctypes.data
        self._fill_tau_source(io, keep, n_col, tau, tau_id, T_tau, q, host=True)
        out = {} if out is None else out
        shapes = {'up': (n_col, L, B), 'down': (n_col, L, B), 'net': (n_col, L),
                  'heating': (n_col, L - 1), 'T_interface': (n_col, L)}
        for name in outputs:
            if name not in out:
                out[name] = np.zeros(shapes[name])
            a = out[name]
            if a.shape != shapes[name] or a.dtype != np.float64 or not a.flags.c_contiguous:
                raise ValueError('out[%r] must be a C-contiguous float64 array of shape %s' % (name, shapes[name]))
            setattr(io, name, a.ctypes.data)
        check(self._lib.cmb200_flux_host(self._h, C.byref(io)))
        return out

    def _fill_tau_source(self, io, keep, n_col, tau, tau_id, T_tau, q, host):
        L, N = self.L, self.N
        if tau is not None:
            if host:
                tau = as_f64(tau)
                keep.append(tau)
                ptr = tau.ctypes.data
            else:
                ptr = tau.data_ptr()
            shape = tuple(tau.shape)
            if len(shape) == 2:
                shape = (1,) + shape
            if shape[1:] != (L, N):
                raise ValueError('tau must have shape [L, N] or [n_tau, L, N] = [.., %d, %d], got %s'
                                 % (L, N, tuple(tau.shape)))
            io.tau, io.n_tau = ptr, shape[0]
            if tau_id is not None:
                if host:
                    tau_id = np.ascontiguousarray(tau_id, dtype=np.int32).reshape(-1)
                    if tau_id.size != n_col:
                        raise ValueError('tau_id must have one entry per column')
                    keep.append(tau_id)
                    io.tau_id = tau_id.ctypes.data
                else:
                    io.tau_id = tau_id.data_ptr()
            return
        if q is None:
            raise ValueError('either tau or q (with tables set on the plan) is required')
        if host:
            q = as_f64(q)
            keep.append(q)
            qptr = q.ctypes.data
        else:
            qptr = q.data_ptr()
        qs = tuple(q.shape)
        if qs == (self.M, L):
            io.q_per_column = 0
        elif qs == (self.M, n_col, L):
            io.q_per_column = 1
        else:
            raise ValueError('q must have shape [M, L] or [M, n_col, L] with M=%d, got %s' % (self.M, qs))
        io.q = qptr
        if T_tau is not None:
            if host:
                T_tau = as_f64(T_tau)
                if T_tau.ndim == 1:
                    T_tau = np.ascontiguousarray(np.broadcast_to(T_tau, (n_col, L)))
                keep.append(T_tau)
                io.T_tau = T_tau.ctypes.data
            else:
                io.T_tau = T_tau.data_ptr()
            if tuple(T_tau.shape) != (n_col, L):
                raise ValueError('T_tau must have shape [n_col, L]')

    # -- device (torch) path ----------------------------------------------------------------
    def flux_device(self, T, T_g, tau=None, tau_id=None, T_tau=None, q=None, albedo=0.3,
                    outputs=('net', 'heating'), out=None):
        """As ``flux`` with torch CUDA float64 tensors (int32 for tau_id); runs on the current
        stream and returns torch tensors without synchronising."""
        import torch
        L, B = self.L, self.B
        dev = torch.device('cuda', self.device)

        def chk(t, dtype=torch.float64):
            if t.device != dev or t.dtype != dtype or not t.is_contiguous():
                raise ValueError('expected a contiguous %s tensor on %s' % (dtype, dev))
            return t
        T = chk(T)
        n_col = T.shape[0]
        if tuple(T.shape) != (n_col, L - 1):
            raise ValueError('T must have shape [n_col, %d]' % (L - 1))
        T_g = chk(T_g)
        io = FluxIO()
        io.n_col = n_col
        io.T, io.T_g = T.data_ptr(), T_g.data_ptr()
        if torch.is_tensor(albedo):
            io.albedo = chk(albedo).data_ptr()
        else:
            io.albedo, io.albedo_scalar = None, float(albedo)
        if tau is not None:
            chk(tau)
            if tau_id is not None:
                chk(tau_id, torch.int32)
        else:
            chk(q)
            if T_tau is not None:
                chk(T_tau)
        self._fill_tau_source(io, [], n_col, tau, tau_id, T_tau, q, host=False)
        res = {} if out is None else out
        shapes = {'up': (n_col, L, B), 'down': (n_col, L, B), 'net': (n_col, L),
                  'heating': (n_col, L - 1), 'T_interface': (n_col, L)}
        for name in outputs:
            if name not in res:
                res[name] = torch.zeros(shapes[name], dtype=torch.float64, device=dev)
            setattr(io, name, chk(res[name]).data_ptr())
        stream = torch.cuda.current_stream(dev).cuda_stream
        check(self._lib.cmb200_flux(self._h, C.byref(io), C.c_void_p(stream)))
        return res

    def flux_sharded(self, T, T_g, T_tau, q, albedo, out, group=None, n_chunks=16):
        """Host buffers in, host buffers out, on a plan restricted by ``shard_by_wavenumber``
        (SURVEY.md 8e): every rank streams the SAME columns through its share of the spectrum and
        the partial net flux / heating rate are added across ranks on the device, chunk by chunk
        (NCCL all-reduce, the path's one exchange step), before they are copied out -- so after
        the call ``out['net']`` [n_col, L] and ``out['heating']`` [n_col, L-1] hold the complete
        result on every rank.  T [n_col, L-1], T_g [n_col], T_tau [n_col, L], q [M, n_col, L] or
        [M, L], albedo [n_col]: float64 NumPy arrays, page-locked if the copies are to overlap
        the kernels (three streams: copy in / kernel + all-reduce / copy out)."""
        import torch
        import torch.distributed as dist
        L, M = self.L, self.M
        dev = torch.device('cuda', self.device)
        n_col = T.shape[0]
        per_col_q = q.ndim == 3
        step = max(1, -(-n_col // n_chunks))
        step = min(n_col, -(-step // 1184) * 1184)                 # whole waves of 148 CTAs x 8 warps
        st = getattr(self, '_sharded_state', None)
        if st is None or st['step'] != step or st['per_col_q'] != per_col_q:
            def buf(*shape):
                return [torch.empty(shape, dtype=torch.float64, device=dev) for _ in range(2)]
            st = {'step': step, 'per_col_q': per_col_q, 'T': buf(step, L - 1), 'T_g': buf(step),
                  'T_tau': buf(step, L), 'albedo': buf(step),
                  'q': buf(M, step, L) if per_col_q else None, 'res': buf(step * (2 * L - 1)),
                  'streams': [torch.cuda.Stream(dev) for _ in range(3)],
                  'in_free': [None, None], 'res_free': [None, None]}
            self._sharded_state = st
        s_in, s_comp, s_out = st['streams']
        hT, hTg, hTt, hq, hal = (torch.from_numpy(x) for x in (T, T_g, T_tau, q, albedo))
        h_net, h_heat = torch.from_numpy(out['net']), torch.from_numpy(out['heating'])
        q_dev = None if per_col_q else hq.to(dev)
        last_out = None
        for i, c0 in enumerate(range(0, n_col, step)):
            c1, n, b = min(n_col, c0 + step), min(n_col, c0 + step) - c0, i & 1
            with torch.cuda.stream(s_in):
                if st['in_free'][b] is not None:
                    s_in.wait_event(st['in_free'][b])                # the kernel that read this buffer is done
                st['T'][b][:n].copy_(hT[c0:c1], non_blocking=True)
                st['T_g'][b][:n].copy_(hTg[c0:c1], non_blocking=True)
                st['T_tau'][b][:n].copy_(hTt[c0:c1], non_blocking=True)
                st['albedo'][b][:n].copy_(hal[c0:c1], non_blocking=True)
                if per_col_q:
                    qb = st['q'][b].view(-1)[:M * n * L].view(M, n, L)
                    for m in range(M):
                        qb[m].copy_(hq[m, c0:c1], non_blocking=True)
                ready = s_in.record_event()
            with torch.cuda.stream(s_comp):
                s_comp.wait_event(ready)
                if st['res_free'][b] is not None:
                    s_comp.wait_event(st['res_free'][b])              # its previous content has been copied out
                res = st['res'][b][:n * (2 * L - 1)]
                views = {'net': res[:n * L].view(n, L), 'heating': res[n * L:].view(n, L - 1)}
                self.flux_device(st['T'][b][:n], st['T_g'][b][:n], T_tau=st['T_tau'][b][:n],
                                 q=st['q'][b].view(-1)[:M * n * L].view(M, n, L) if per_col_q else q_dev,
                                 albedo=st['albedo'][b][:n], outputs=('net', 'heating'), out=views)
                st['in_free'][b] = s_comp.record_event()
                if dist.is_initialized() and dist.get_world_size(group) > 1:
                    dist.all_reduce(res, group=group)
                done = s_comp.record_event()
            with torch.cuda.stream(s_out):
                s_out.wait_event(done)
                h_net[c0:c1].copy_(views['net'], non_blocking=True)
                h_heat[c0:c1].copy_(views['heating'], non_blocking=True)
                st['res_free'][b] = last_out = s_out.record_event()
        if last_out is not None:
            last_out.synchronize()
        st['in_free'], st['res_free'] = [None, None], [None, None]
        return out

    def optical_depth_device(self, q, T_tau=None, T_g=None, n_col=None):
        """tau [n_col, L, N] (torch, float64) from the plan's tables (optical_depth,
        real_gas.py:86-127, on the plan grid)."""
        import torch
        dev = torch.device('cuda', self.device)
        if T_tau is not None:
            n_col = T_tau.shape[0]
        elif T_g is not None:
            n_col = T_g.shape[0]
        per_col = 1 if q.dim() == 3 else 0
        tau = torch.empty((n_col, self.L, self.N), dtype=torch.float64, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        check(self._lib.cmb200_plan_optical_depth(
            self._h, n_col, T_tau.data_ptr() if T_tau is not None else None,
            T_g.data_ptr() if T_g is not None else None, q.data_ptr(), per_col, tau.data_ptr(),
            C.c_void_p(stream)))
        return tau

    def optical_depth_noalloc(self, q, tau, T_tau=None, T_g=None):
        """``optical_depth_device`` for callers that must not allocate (CUDA graph capture, serving loops):
        the caller passes the output ``tau`` [n_col, L, N] (contiguous float64 CUDA tensor) and
        cmb200_plan_optical_depth_noalloc fills it using a fixed 1 MiB device array as scratch, so
        M * n_col * L <= 1,048,576 (ValueError beyond).  Calls that overlap on the GPU race on that
        array.  No counterpart in the reference."""
        import torch
        n_col = tau.shape[0]
        if tau.shape != (n_col, self.L, self.N) or tau.dtype != torch.float64 or not tau.is_contiguous():
            raise ValueError('tau must be a contiguous float64 tensor [n_col, %d, %d]' % (self.L, self.N))
        per_col = 1 if q.dim() == 3 else 0
        stream = torch.cuda.current_stream(tau.device).cuda_stream
        check(self._lib.cmb200_plan_optical_depth_noalloc(
            self._h, n_col, T_tau.data_ptr() if T_tau is not None else None,
            T_g.data_ptr() if T_g is not None else None, q.data_ptr(), per_col, tau.data_ptr(),
            C.c_void_p(stream)))
        return tau
