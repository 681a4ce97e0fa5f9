"""BatchEngine -- host-side owner of one C-ABI handle (one GPU, one population of episodes).

PyTorch is used for device memory, streams and host<->device copies only; every simulated minute is
computed by the CUDA library behind include/simglucose_b200.h.  All arrays that cross into the
library are structure-of-arrays device tensors of shape [rows, n_episodes].
"""
import ctypes as C

import numpy as np
import torch

from . import _capi
from . import params as P


def _stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


class BatchEngine:
    """n_episodes independent closed loops resident on one GPU.

    Mirrors, for a whole population at once, what the reference does per patient with
    T1DSimEnv (simulation/env.py:36-180) driven by SimObj.simulate (simulation/sim_engine.py:29-39).
    """

    def __init__(self, capacity, device=0, carry_h=False, fma=False, sensor='Dexcom', pump='Insulet', fp32=False,
                 coherence=0, adaptive_sort=False):
        """coherence: env-steps between two sorts of the launch order (set_coherence); 0 = off.  adaptive_sort: sort by
        what every episode just cost instead of by its next meal event (set_coherence)."""
        if not torch.cuda.is_available():
            raise _capi.SgError('simglucose_b200 needs a CUDA device (there is no CPU fallback)')
        self.lib = _capi.load()
        self.device = torch.device('cuda', device)
        self.capacity = int(capacity)
        self.flags = ((_capi.FLAG_CARRY_H if carry_h else 0) | (_capi.FLAG_FMA if fma else 0) |
                      (_capi.FLAG_FP32 if fp32 else 0))
        h = C.c_void_p()
        _capi.check(self.lib.sg_create(self.device.index, self.capacity, self.flags, C.byref(h)))
        self._h = h
        self.n = 0
        self._keep = {}
        self.set_hardware(sensor, pump)
        tbl, x0 = P.cohort_table()
        self.load_table(tbl, x0)
        self.step_cap = 0
        self.coherence = 0
        self.adaptive_sort = False
        if coherence:
            self.set_coherence(coherence, adaptive=adaptive_sort)

    # -- configuration ---------------------------------------------------------------------------
    def set_hardware(self, sensor='Dexcom', pump='Insulet'):
        s = P.sensor_params(sensor) if isinstance(sensor, str) else sensor
        p = P.pump_params(pump) if isinstance(pump, str) else pump
        hw = _capi.Hardware()
        hw.sample_time = int(s['sample_time'])
        hw.sensor_min, hw.sensor_max = float(s['min']), float(s['max'])
        hw.pump[:] = [p['min_basal'], p['max_basal'], p['inc_basal'], p['min_bolus'], p['max_bolus'], p['inc_bolus']]
        _capi.check(self.lib.sg_set_hardware(self._h, C.byref(hw)))
        self.sample_time = hw.sample_time
        self.sensor, self.pump = s, p

    def load_table(self, table, x0):
        table = np.ascontiguousarray(table, dtype=np.float64)
        x0 = np.ascontiguousarray(x0, dtype=np.float64)
        assert table.ndim == 2 and table.shape[1] == _capi.NPARAM and x0.shape == (table.shape[0], _capi.NX)
        dp = C.POINTER(C.c_double)
        _capi.check(self.lib.sg_load_params(self._h, table.ctypes.data_as(dp), x0.ctypes.data_as(dp), table.shape[0]))
        self.n_rows = table.shape[0]

    # -- helpers -----------------------------------------------------------------------------------
    def _dev(self, a, dtype, rows=None):
        """-> contiguous device tensor [rows, n]; accepts numpy / torch, host / device."""
        if a is None:
            return None
        t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))
        t = t.to(device=self.device, dtype=dtype, non_blocking=True)
        if t.ndim == 1:
            t = t.unsqueeze(0)
        t = t.contiguous()
        assert t.shape[-1] == self.n, 'last dimension must be n_episodes (structure-of-arrays)'
        if rows is not None:
            assert t.shape[0] == rows, (t.shape, rows)
        return t

    def empty(self, *rows, dtype=torch.float64):
        return torch.empty(*rows, self.n, dtype=dtype, device=self.device)

    # -- episode life cycle --------------------------------------------------------------------
    def reset(self, patient_row, noise, meal_minute=None, meal_grams=None, x0=None, params=None, row0=True):
        """T1DSimEnv.reset for the whole population.  All arrays SoA: [rows, E].

        patient_row [E] int32; noise [n_noise, E]; meal_minute/meal_grams [M, E]; x0 [13, E] (None: table);
        params [32, E] (None: table).  Returns history row 0 as a [NTRACE, E] tensor (or None)."""
        pr = patient_row if isinstance(patient_row, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(patient_row))
        self.n = int(pr.numel())
        if self.n < 1 or self.n > self.capacity:
            raise ValueError('n_episodes %d outside 1..capacity %d' % (self.n, self.capacity))
        pr = pr.to(device=self.device, dtype=torch.int32).reshape(self.n).contiguous()
        nz = self._dev(noise, torch.float64)
        mm = self._dev(meal_minute, torch.int32)
        mg = self._dev(meal_grams, torch.float64)
        max_meals = 0 if mm is None else mm.shape[0]
        if mm is not None:
            assert mg is not None and mg.shape == mm.shape
        x0t = self._dev(x0, torch.float64, _capi.NX)
        pt = self._dev(params, torch.float64, _capi.NPARAM)
        r0 = self.empty(_capi.NTRACE) if row0 else None
        # the library borrows these until the next reset
        self._keep = dict(pr=pr, nz=nz, mm=mm, mg=mg, x0=x0t, pt=pt)
        _capi.check(self.lib.sg_reset(self._h, self.n, _ptr(pr), _ptr(pt), _ptr(x0t), _ptr(mm), _ptr(mg), max_meals,
                                      _ptr(nz), nz.shape[0], _ptr(r0), _stream_ptr()))
        self.steps_done = 0
        return r0

    def step(self, basal, bolus, out=None):
        """T1DSimEnv.step for every episode: one env-step with Action(basal[e], bolus[e])."""
        b = self._dev(basal, torch.float64, 1)
        o = self._dev(bolus, torch.float64, 1)
        if out is None:
            out = self.empty(_capi.NTRACE)
        _capi.check(self.lib.sg_step(self._h, _ptr(b), _ptr(o), _ptr(out), _stream_ptr()))
        self.steps_done += 1
        return out

    def step_meals(self, basal, bolus, meals, out=None):
        """T1DSimEnv.step with this step's announced meals supplied by the caller: meals [sample_time, E] =
        scenario.get_action(time).meal for each minute of the step (env.py:50) -- the seam for arbitrary
        Scenario subclasses evaluated on the host."""
        b = self._dev(basal, torch.float64, 1)
        o = self._dev(bolus, torch.float64, 1)
        m = self._dev(meals, torch.float64, self.sample_time)
        if out is None:
            out = self.empty(_capi.NTRACE)
        _capi.check(self.lib.sg_step_meals(self._h, _ptr(b), _ptr(o), _ptr(m), _ptr(out), _stream_ptr()))
        self.steps_done += 1
        return out

    def patient_step(self, cho, insulin, out=None):
        """T1DPatient.step for every episode: one minute with Action(CHO = grams announced, insulin U/min).
        Returns [NPOUT, E]: x[13], Gsub, to_eat, idid, dopri5 steps, planned_meal, _last_Qsto,
        _last_foodtaken, is_eating."""
        c = self._dev(cho, torch.float64, 1)
        i = self._dev(insulin, torch.float64, 1)
        if out is None:
            out = self.empty(_capi.NPOUT)
        _capi.check(self.lib.sg_patient_step(self._h, _ptr(c), _ptr(i), _ptr(out), _stream_ptr()))
        return out

    def set_step_cap(self, cap):
        """Heavy-minute compaction (sg_set_step_cap): minutes needing more than `cap` DOPRI5 steps are finished by a
        second launch over the compacted list.  For populations with independent meal schedules; 0 = off."""
        _capi.check(self.lib.sg_set_step_cap(self._h, int(cap)))
        self.step_cap = int(cap)

    def set_coherence(self, segment_steps, adaptive=False):
        """Schedule-coherent launch order (sg_set_coherence): every `segment_steps` env-steps the episodes are sorted
        by the minute of their next meal event, so that the lanes of a warp meet the expensive minutes together.
        For populations whose episodes have meal schedules of their own; per-episode results do not change.  0 = off.
        adaptive (sg_set_sort_mode 1): sort by the DOPRI5 steps every episode took since the previous sort and by the
        meals announced inside the segment -- for orders that are re-sorted every few env-steps (lockstep stepping of
        environments that restart one by one)."""
        _capi.check(self.lib.sg_set_coherence(self._h, int(segment_steps)))
        _capi.check(self.lib.sg_set_sort_mode(self._h, 1 if adaptive else 0))
        self.coherence = int(segment_steps)
        self.adaptive_sort = bool(adaptive)

    def launch_order(self):
        """The current sorted launch order [E] int32 (thread i owns episode perm[i])."""
        out = torch.empty(self.n, dtype=torch.int32, device=self.device)
        _capi.check(self.lib.sg_get_launch_order(self._h, _ptr(out), _stream_ptr()))
        return out

    def reset_masked(self, mask, x0=None, row0=None):
        """T1DSimEnv.reset for the episodes with mask != 0 only (sg_reset_masked).  The caller has already rewritten
        the masked columns of the noise / meal buffers handed to reset() (they are borrowed, not copied: in-place
        writes to those tensors are what the kernels read).  x0 [13, E]: post-reset states (None: table x0).
        row0 [NTRACE, E]: receives history row 0 of the masked episodes (other columns untouched)."""
        m = self._dev(mask, torch.int32, 1)
        x0t = self._dev(x0, torch.float64, _capi.NX)
        self._keep.update(mask=m, x0m=x0t)
        _capi.check(self.lib.sg_reset_masked(self._h, _ptr(m), _ptr(x0t), _ptr(row0), _stream_ptr()))
        return row0

    def set_t0(self, t0):
        """patient.t at the next reset (T1DPatient(t0=...), t1dpatient.py:27,277)."""
        _capi.check(self.lib.sg_set_t0(self._h, int(t0)))

    def rebind_streams(self, noise, meal_minute=None, meal_grams=None):
        """Swap in longer noise / meal buffers without resetting the episodes."""
        nz = self._dev(noise, torch.float64)
        mm = self._dev(meal_minute, torch.int32)
        mg = self._dev(meal_grams, torch.float64)
        max_meals = 0 if mm is None else mm.shape[0]
        _capi.check(self.lib.sg_rebind_streams(self._h, _ptr(mm), _ptr(mg), max_meals, _ptr(nz), nz.shape[0]))
        torch.cuda.current_stream().synchronize()       # kernels in flight may still read the old buffers
        self._keep.update(nz=nz, mm=mm, mg=mg)

    def run(self, n_steps, controller='bb', ctrl=None, ctrl_rows=None, actions=None, trace=True, counters=False):
        """SimObj.simulate for every episode, n_steps env-steps in one launch.

        controller: 'bb' (ctrl = target, CR, CF, basal -- basal 0: from the episode's own parameter row), 'pid'
        (ctrl = P, I, D, target), 'random' (ctrl = lo, hi, seed, global index of episode 0: basal ~ U(lo, hi) drawn in
        the kernel) or 'actions' (actions [n_steps, 2, E]).  ctrl_rows [4, E] overrides ctrl per episode.
        trace may be a preallocated [n_steps, NTRACE, E] tensor.
        Returns dict(trace [n_steps, NTRACE, E] | None, counters [n_steps*st, 2, E] | None)."""
        kind = {'actions': _capi.CTRL_ACTIONS, 'bb': _capi.CTRL_BB, 'pid': _capi.CTRL_PID,
                'random': _capi.CTRL_RANDOM}[controller]
        cbuf = None
        if ctrl is not None:
            c = list(ctrl) + [0.0] * (_capi.NCTRL - len(ctrl))
            cbuf = (C.c_double * _capi.NCTRL)(*c)
        cr = self._dev(ctrl_rows, torch.float64, _capi.NCTRL)
        act = None
        if actions is not None:
            act = actions if isinstance(actions, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(actions))
            act = act.to(device=self.device, dtype=torch.float64).contiguous()
            assert tuple(act.shape) == (n_steps, 2, self.n)
        if isinstance(trace, torch.Tensor):
            tr = trace
            assert tr.is_contiguous() and tr.numel() == n_steps * _capi.NTRACE * self.n and tr.dtype == torch.float64
        else:
            tr = torch.empty(n_steps, _capi.NTRACE, self.n, dtype=torch.float64, device=self.device) if trace else None
        cn = (torch.empty(n_steps * self.sample_time, 2, self.n, dtype=torch.int32, device=self.device)
              if counters else None)
        _capi.check(self.lib.sg_run(self._h, kind, cbuf, _ptr(cr), _ptr(act), int(n_steps), _ptr(tr), _ptr(cn),
                                    _stream_ptr()))
        self.steps_done += n_steps
        return dict(trace=tr, counters=cn)

    # -- readers -----------------------------------------------------------------------------------
    def state(self):
        out = self.empty(_capi.NX)
        _capi.check(self.lib.sg_get_state(self._h, _ptr(out), _stream_ptr()))
        return out

    def episode_state(self):
        """The whole resident state: (d [ND, E] float64, i [NI, E] int32); rows: include/simglucose_b200.h
        sg_state_d / sg_state_i."""
        d = self.empty(_capi.ND)
        i = self.empty(_capi.NI, dtype=torch.int32)
        _capi.check(self.lib.sg_get_episode_state(self._h, _ptr(d), _ptr(i), _stream_ptr()))
        return d, i

    def set_episode_state(self, d=None, i=None, steps_done=None):
        """Overwrite the resident state; steps_done = env-steps those episodes have already taken (keeps the
        noise-stream length checks valid)."""
        dd = self._dev(d, torch.float64, _capi.ND)
        ii = self._dev(i, torch.int32, _capi.NI)
        sd = self.steps_done if steps_done is None else int(steps_done)
        _capi.check(self.lib.sg_set_episode_state(self._h, _ptr(dd), _ptr(ii), sd, _stream_ptr()))
        self.steps_done = sd
        torch.cuda.current_stream().synchronize()

    def model(self, x, cho, insulin, last_qsto, last_food, patient_row=None, params=None):
        """T1DPatient.model for n independent evaluations: x [13, n]; cho, insulin, last_qsto, last_food [n];
        params [32, n] or patient_row [n] into the loaded table.  Returns dxdt [13, n]."""
        def dev(a, dtype=torch.float64):
            t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))
            return t.to(device=self.device, dtype=dtype).contiguous()
        xt = dev(x)
        n = xt.shape[-1]
        assert xt.shape == (_capi.NX, n)
        args = [dev(np.broadcast_to(np.asarray(a, dtype=np.float64), (n,)).copy() if not isinstance(a, torch.Tensor) else a)
                for a in (cho, insulin, last_qsto, last_food)]
        pr = dev(patient_row, torch.int32) if patient_row is not None else None
        pt = dev(params) if params is not None else None
        if pt is not None:
            assert pt.shape == (_capi.NPARAM, n)
        out = torch.empty(_capi.NX, n, dtype=torch.float64, device=self.device)
        _capi.check(self.lib.sg_model(self._h, n, _ptr(xt), _ptr(args[0]), _ptr(args[1]), _ptr(args[2]), _ptr(args[3]),
                                      _ptr(pr), _ptr(pt), _ptr(out), _stream_ptr()))
        return out

    def stats(self):
        out = self.empty(_capi.NSTAT)
        _capi.check(self.lib.sg_get_stats(self._h, _ptr(out), _stream_ptr()))
        return out

    def status(self):
        out = torch.empty(self.n, dtype=torch.int32, device=self.device)
        _capi.check(self.lib.sg_get_status(self._h, _ptr(out), _stream_ptr()))
        return out

    def reduce_stats(self, out=None):
        """Population aggregate [NSTAT] (device tensor): sums, BG extrema, #episodes done."""
        if out is None:
            out = torch.empty(_capi.NSTAT, dtype=torch.float64, device=self.device)
        _capi.check(self.lib.sg_reduce_stats(self._h, _ptr(out), _stream_ptr()))
        return out

    @property
    def launches(self):
        return int(self.lib.sg_launch_count(self._h))

    def close(self):
        if getattr(self, '_h', None) is not None and self._h:
            self.lib.sg_destroy(self._h)
            self._h = None
        self._keep = {}

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ShardRunner:
    """batch_sim for a population that lives in HOST memory as shards of episodes (sim_engine.py:65-76 hands the
    pathos pool one SimObj at a time; here a shard is the unit).  Shard k+1 is uploaded on one CUDA stream while
    shard k is simulated on another: `slots` engines, one stream each, so the host->device copies (8 bytes of
    sensor noise per env-step per episode dominate) hide behind the episode kernel and the tail of one shard's
    launch overlaps the head of the next.  Engines and streams live as long as the runner, so that repeated
    populations reuse their device buffers (PyTorch caches device blocks per stream)."""

    def __init__(self, capacity, device=0, slots=2, **engine_kw):
        self.device = torch.device('cuda', device)
        self.capacity = int(capacity)
        self.engines = [BatchEngine(self.capacity, device=device, **engine_kw) for _ in range(max(1, int(slots)))]
        self.streams = [torch.cuda.Stream(device=self.device) for _ in self.engines]
        self._trace_dev = self._trace_sel = None
        self._trace_shape = (0, 0)

    def run(self, shards, n_steps, controller='bb', ctrl=None, stats_out=None, traces_out=None,
            trace_cols=('BG', 'CGM', 'CHO', 'insulin')):
        """shards: iterable of dicts with host tensors / arrays, structure-of-arrays [rows, E_k] (pinned memory makes
          the copies asynchronous): patient_row [E_k], noise [n_steps + 2, E_k], meal_minute / meal_grams [M, E_k]
          and optionally ctrl_rows [4, E_k], x0 [13, E_k], params [32, E_k].
        stats_out: optional list of pinned [NSTAT, E_k] tensors that receive the per-episode statistics.
        traces_out: optional list of pinned [n_steps, len(trace_cols), E_k] tensors that receive the trajectories (the
          columns of the reference's result DataFrame, sim_engine.py:44-49 / env.py:169-180, one row per env-step; the
          reset row is what reset(row0=...) returns); the device -> host copy of shard k overlaps the kernel of k + 1.
        Returns (stats_out, aggregate) -- aggregate is the population statistics vector [NSTAT] (numpy): sums, BG
        extrema (rows 2, 3), episodes that reported done (row 13).
        (Measured and dropped: uploading the FIRST shard's noise in slices of env-steps on a copy stream, its run cut
        likewise, so that the kernel starts after a third of the upload -- 4.6 % slower end to end than this plain
        two-stream pipeline, profiles/r02_i_*.)"""
        shards = list(shards)
        cols = [_capi.TRACE_COLS.index(c) for c in trace_cols]
        if traces_out is not None:
            # staging buffers are kept between calls and only ever grow; the column selection belongs to THIS call
            need = (int(n_steps), len(cols))
            if self._trace_dev is None or self._trace_shape[0] < need[0] or self._trace_shape[1] < need[1]:
                have = self._trace_shape if self._trace_dev is not None else (0, 0)
                self._trace_shape = (max(need[0], have[0]), max(need[1], have[1]))
                self._trace_dev = [torch.empty(self._trace_shape[0], _capi.NTRACE, self.capacity, dtype=torch.float64,
                                               device=self.device) for _ in self.engines]
                self._trace_sel = [torch.empty(self._trace_shape[0], self._trace_shape[1], self.capacity,
                                               dtype=torch.float64, device=self.device) for _ in self.engines]
            if getattr(self, '_cols_key', None) != tuple(cols):
                self._cols = torch.tensor(cols, dtype=torch.int64, device=self.device)
                self._cols_key = tuple(cols)
        if stats_out is None:
            stats_out = [torch.empty(_capi.NSTAT, int(torch.as_tensor(s['patient_row']).numel()),
                                     dtype=torch.float64).pin_memory() for s in shards]
        aggs = torch.empty(len(shards), _capi.NSTAT, dtype=torch.float64).pin_memory()
        here = torch.cuda.current_stream(self.device)
        for st in self.streams:
            st.wait_stream(here)
        n = len(self.engines)
        for k, sh in enumerate(shards):
            eng, st = self.engines[k % n], self.streams[k % n]
            with torch.cuda.stream(st):
                eng.reset(sh['patient_row'], sh['noise'], sh.get('meal_minute'), sh.get('meal_grams'),
                          x0=sh.get('x0'), params=sh.get('params'), row0=False)
                if traces_out is None:
                    eng.run(n_steps, controller=controller, ctrl=ctrl, ctrl_rows=sh.get('ctrl_rows'), trace=False)
                else:
                    ek = eng.n
                    tr = self._trace_dev[k % n].view(-1)[:n_steps * _capi.NTRACE * ek].view(n_steps, _capi.NTRACE, ek)
                    eng.run(n_steps, controller=controller, ctrl=ctrl, ctrl_rows=sh.get('ctrl_rows'), trace=tr)
                    sel = self._trace_sel[k % n].view(-1)[:n_steps * len(cols) * ek].view(n_steps, len(cols), ek)
                    torch.index_select(tr, 1, self._cols, out=sel)
                    traces_out[k].copy_(sel, non_blocking=True)
                stats_out[k].copy_(eng.stats(), non_blocking=True)
                aggs[k].copy_(eng.reduce_stats(), non_blocking=True)
        for st in self.streams:
            here.wait_stream(st)
            st.synchronize()
        a = aggs.numpy()
        agg = a.sum(axis=0)
        agg[2], agg[3] = a[:, 2].min(), a[:, 3].max()
        return stats_out, agg

    def close(self):
        for eng in self.engines:
            eng.close()
        self.engines = []
        self._trace_dev = self._trace_sel = None


def run_shards(shards, n_steps, controller='bb', ctrl=None, device=0, slots=2, stats_out=None, **engine_kw):
    """One-shot ShardRunner: see ShardRunner.run for the shard format and the return value."""
    shards = list(shards)
    cap = max(int(torch.as_tensor(s['patient_row']).numel()) for s in shards)
    runner = ShardRunner(cap, device=device, slots=min(int(slots), len(shards)), **engine_kw)
    try:
        return runner.run(shards, n_steps, controller=controller, ctrl=ctrl, stats_out=stats_out)
    finally:
        runner.close()
