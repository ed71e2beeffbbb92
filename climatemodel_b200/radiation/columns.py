"""Many independent columns stepped on the device: the flux evaluation, the adaptive time step
of ``Atmosphere.update_temp`` (base.py:166-246) and, optionally, the dry convective adjustment
(convective_adjustment.py:7-135) run back to back on the GPU, every column with its own dt,
without a host round trip per step -- the batched form of ``RealGas.take_time_step``
(BASELINE configs[2]).

    cols = ColumnStepper(plan, T, T_g, q=q, T_tau=T_tau, albedo=albedo)     # torch CUDA tensors
    for _ in range(n):
        delta_net_flux = cols.step()
    cols.T, cols.t                                                          # [n_col, L-1], [n_col]
"""
import ctypes as C

import numpy as np

from .. import _lib
from .._lib import check


class ColumnStepper:
    def __init__(self, plan, T, T_g, q=None, T_tau=None, tau=None, tau_id=None, albedo=0.3, temp_change=1,
                 delta_temp_change=0.01, level_order='top_first'):
        """plan: FluxPlan; T [n_col, L-1], T_g [n_col] and the optical-depth source (q, T_tau or
        tau, tau_id) as for ``FluxPlan.flux_device`` (torch float64 CUDA tensors).  T is advanced in
        place.  Level index 0 is the top of the atmosphere, as in ``FluxPlan``.

        level_order: the level numbering of the model each column stands for -- it decides which
        interface's |net flux| gates a layer (base.py:171), arg-max ties (base.py:209) and the
        numbering of MaxTendInd.  'top_first' (default) = ``RealGas`` (p_interface ascending,
        real_gas.py:281); 'surface_first' = an ``Atmosphere`` whose arrays start at the surface."""
        if level_order not in ('top_first', 'surface_first'):
            raise ValueError("level_order must be 'top_first' or 'surface_first'")
        import torch
        self.plan, self.T, self.T_g = plan, T, T_g
        self.source = dict(q=q, T_tau=T_tau, tau=tau, tau_id=tau_id, albedo=albedo)
        self.n_col = T.shape[0]
        self.t = torch.zeros(self.n_col, dtype=torch.float64, device=T.device)
        self.delta_net_flux = torch.empty(self.n_col, dtype=torch.float64, device=T.device)
        self.out = {}
        self._lib = _lib.load()
        self._h = C.c_void_p()
        check(self._lib.cmb200_stepper_create(C.byref(self._h), plan._h, self.n_col, float(temp_change),
                                              float(delta_temp_change)))
        self.level_order = level_order
        check(self._lib.cmb200_stepper_set_level_order(self._h, int(level_order == 'top_first')))
        # cell centres in the reference's order (surface first) and (p / p_ref)^alpha as NumPy evaluates it
        from .convective_adjustment import g, c_p_dry, p_surface_earth, R_specific
        P_ref = np.asarray(plan.P, dtype=float)[::-1]
        self.p_centre = np.ascontiguousarray((P_ref[:-1] + P_ref[1:]) / 2)
        alpha = R_specific * (g / c_p_dry) / g
        self.p_power = np.ascontiguousarray((self.p_centre / p_surface_earth) ** alpha)
        check(self._lib.cmb200_stepper_set_convection(self._h, _lib.dptr(self.p_centre), _lib.dptr(self.p_power)))

    def __del__(self):
        if getattr(self, '_h', None) is not None and self._h.value:
            self._lib.cmb200_stepper_destroy(self._h)
            self._h = None

    def reset(self):
        """Back to t = 0 with the initial time-step state (T is left as it is)."""
        check(self._lib.cmb200_stepper_reset(self._h))
        self.t.zero_()

    def step(self, changing_tau=False, convective_adjust=False, net_flux_thresh=1e-7, net_flux_percentile=95,
             conv_thresh=1e-5, conv_t_multiplier=5):
        """One adaptive step of every column (arguments as ``Atmosphere.update_temp``); returns
        delta_net_flux [n_col] (device tensor)."""
        import torch
        self.out = self.plan.flux_device(self.T, self.T_g, outputs=('net',), out=self.out, **self.source)
        stream = torch.cuda.current_stream(self.T.device).cuda_stream
        head = (self._h, self.out['net'].data_ptr(), self.T.data_ptr(), self.t.data_ptr(),
                self.delta_net_flux.data_ptr(), float(net_flux_thresh), float(net_flux_percentile),
                int(bool(changing_tau)))
        if convective_adjust:
            check(self._lib.cmb200_stepper_update_convective(*head, float(conv_thresh), float(conv_t_multiplier),
                                                             C.c_void_p(stream)))
        else:
            check(self._lib.cmb200_stepper_update(*head, C.c_void_p(stream)))
        return self.delta_net_flux

    def evolve_to_equilibrium(self, flux_thresh=1e-3, convective_adjust=False, t_end=4.0, conv_thresh=1e-5,
                              conv_t_multiplier=5, max_steps=1000000, check_every=16, use_graph=False):
        """``Atmosphere.evolve_to_equilibrium`` (base.py:266-335) for every column at once: all columns
        step together, and a column stops -- is frozen on the device, its T, t and step state left as
        they are -- as soon as ITS equilibrium test holds: max|net flux| or delta_net_flux below its
        threshold (base.py:248-264; the threshold drops to 0.99 delta_net_flux of the column's second
        step, base.py:315-316) or more than t_end years of model time.  The tests run on the device
        with no host round trip; the host looks at the "all done" flag every check_every steps.

        use_graph: after the first two steps (which differ: the threshold update) one whole step --
        flux kernels, adaptive step, convective adjustment, equilibrium test, freeze mask -- is captured
        into a CUDA graph and replayed, so a step costs one launch instead of ~20.  Same kernels on the
        same data in the same order: results are identical to the eager loop.  Off by default: measured
        on the 128-column ensemble of BASELINE configs[4] the eager loop already runs at the pace of the
        flux kernel (0.60 ms per step of which 0.51 ms flux: one column per SM), so the graph only pays
        for its capture (0.246 s eager, 0.238-0.292 s with the graph, 408 steps).

        Returns the number of steps each column took (int64 tensor).  Raises ValueError when a
        temperature goes negative (base.py:319-320) or max_steps is hit."""
        import torch
        dev = self.T.device
        n = self.n_col
        thr = torch.full((n,), float(flux_thresh), dtype=torch.float64, device=dev)
        done = torch.zeros(n, dtype=torch.uint8, device=dev)
        n_steps = torch.zeros(n, dtype=torch.int64, device=dev)

        def years(t):                                   # t_years_days (base.py:23-27): years + days / 365
            full_days = t / (24 * 60 ** 2)
            whole = torch.floor(full_days / 365)
            return whole + (full_days - 365 * whole) / 365
        years0 = years(self.t)

        def one_step(update_threshold=False):           # every tensor it touches lives outside: replayable
            running = done == 0
            dnf = self.step(convective_adjust=convective_adjust, conv_thresh=conv_thresh,
                            conv_t_multiplier=conv_t_multiplier)
            n_steps.add_(running)
            if update_threshold:
                thr.copy_(torch.minimum(thr, 0.99 * dnf))
            eq = (self.out['net'].abs().amax(dim=1) < thr) | (dnf < thr) | (years(self.t) - years0 > t_end)
            done.bitwise_or_((eq & running).to(torch.uint8))

        def host_check(i):
            if float(self.T.min()) < 0:
                raise ValueError('Temperature is below zero')
            if bool(done.all()):
                return True
            if i >= max_steps:
                raise ValueError('no equilibrium after %d steps' % max_steps)
            return False

        check(self._lib.cmb200_stepper_set_frozen(self._h, C.c_void_p(done.data_ptr())))
        graph = None
        try:
            i = 0
            while i < 2 and i < max_steps:              # eager: warm-up of every buffer, and the step that sets thr
                one_step(update_threshold=(i == 1))
                i += 1
            finished = host_check(i)
            if not finished and use_graph:
                try:
                    side = torch.cuda.Stream(device=dev)
                    side.wait_stream(torch.cuda.current_stream(dev))
                    with torch.cuda.stream(side):
                        one_step()                      # on the capture stream once, outside the capture
                        i += 1
                    torch.cuda.current_stream(dev).wait_stream(side)
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph, stream=side):
                        one_step()
                except Exception:                       # capture refused (e.g. an allocation inside): eager loop
                    graph = None
                    torch.cuda.synchronize(dev)
            while not finished:
                for _ in range(check_every):
                    if graph is not None:
                        graph.replay()
                    else:
                        one_step()
                    i += 1
                    if i >= max_steps:
                        break
                finished = host_check(i)
        finally:
            torch.cuda.current_stream(dev).synchronize()
            check(self._lib.cmb200_stepper_set_frozen(self._h, None))
            del graph
        self.graph_replayed = use_graph
        return n_steps

    @property
    def net_flux(self):
        return self.out.get('net')

    def state(self):
        """Host copies of the per-column time-step state: DeltaT, dt, MaxTendInd (numbered in
        ``level_order``), number of frozen levels."""
        n = self.n_col
        dT, dt = np.zeros(n), np.zeros(n)
        ind, rem = np.zeros(n, dtype=np.int32), np.zeros(n, dtype=np.int32)
        check(self._lib.cmb200_stepper_state(self._h, _lib.dptr(dT), _lib.dptr(dt),
                                             ind.ctypes.data_as(C.POINTER(C.c_int32)),
                                             rem.ctypes.data_as(C.POINTER(C.c_int32))))
        return {'DeltaT': dT, 'dt': dt, 'MaxTendInd': ind, 'n_removed': rem}


def convective_adjustment_device(p, T, conv_thresh=1e-5):
    """``convective_adjustment(p, T)`` (convective_adjustment.py:118-135) for many columns on the GPU.
    p [nz] cell-centre pressures in either order; T: torch float64 CUDA tensor [n_col, nz] in the
    same level order as p, adjusted in place.  Returns (T, status[n_col]); status 1 marks columns in
    which an unstable run reached the top level (where the reference raises IndexError)."""
    import torch
    from .convective_adjustment import g, c_p_dry, p_surface_earth, R_specific
    p = np.asarray(p, dtype=float).reshape(-1)
    if T.dim() != 2 or T.shape[1] != p.size or T.dtype != torch.float64 or not T.is_contiguous():
        raise ValueError('T must be a contiguous float64 tensor [n_col, %d]' % p.size)
    surface_first = p[0] > p[1]
    p_ref = np.ascontiguousarray(p if surface_first else p[::-1])
    pw = np.ascontiguousarray((p_ref / p_surface_earth) ** (R_specific * (g / c_p_dry) / g))
    dev = T.device
    work = T.flip(1).contiguous() if surface_first else T          # the kernel's level order: top first
    status = torch.zeros(T.shape[0], dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    p_dev, pw_dev = torch.tensor(p_ref, device=dev), torch.tensor(pw, device=dev)      # alive until the sync below
    check(_lib.load().cmb200_convective_adjust(p.size, T.shape[0], p_dev.data_ptr(), pw_dev.data_ptr(),
                                               float(conv_thresh), work.data_ptr(), None, None, status.data_ptr(),
                                               C.c_void_p(stream)))
    torch.cuda.current_stream(dev).synchronize()
    if surface_first:
        T.copy_(work.flip(1))
    return T, status


# ---- independent columns over several ranks (column sharding, no data-path collective) ------
def column_share(n_all, rank, world):
    """Global indices of the columns rank `rank` of `world` steps: round-robin, so that neighbours
    (in latitude, say), which cost about the same, land on different ranks."""
    return np.arange(rank, n_all, world)


def gather_columns(rows, n_all, rank, world, group=None):
    """rows: torch tensor [len(column_share(n_all, rank, world)), k] of per-column results on this
    rank (any device torch.distributed's backend handles).  Returns the complete [n_all, k] table in
    global column order on every rank: one all-gather of equally padded blocks -- the only
    communication of a column-sharded run, after the stepping is done."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return rows
    pad = -(-n_all // world)
    block = torch.zeros((pad, rows.shape[1]), dtype=rows.dtype, device=rows.device)
    block[:rows.shape[0]] = rows
    blocks = [torch.zeros_like(block) for _ in range(world)]
    dist.all_gather(blocks, block, group=group)
    full = torch.zeros((n_all, rows.shape[1]), dtype=rows.dtype, device=rows.device)
    for r in range(world):
        idx = torch.as_tensor(column_share(n_all, r, world), device=rows.device)
        full[idx] = blocks[r][:idx.numel()]
    return full
