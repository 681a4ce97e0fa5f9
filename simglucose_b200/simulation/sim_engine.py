"""SimObj / sim / batch_sim -- the surface of simglucose/simulation/sim_engine.py:13-76.

The reference runs one Python loop per patient and spreads patients over a pathos process pool.  Here
batch_sim gathers every simulation whose plant is this package's (T1DPatient, CGMSensor, InsulinPump) and
whose controller is a BBController or PIDController into ONE population on the GPU and runs all their
episodes in a single launch of the episode kernel (controller fused, sg_run); the per-env histories are then
filled from the trace, so results(), save_results() and the env objects look exactly as after the
reference's loop.  Anything else (a custom Controller subclass, foreign plant objects) runs the reference's
loop, each env-step still one GPU launch.
"""
import logging
import os
import time
from datetime import timedelta

import numpy as np
import torch

from .. import streams
from ..engine import BatchEngine

logger = logging.getLogger(__name__)

pathos = False     # kept for scripts that inspect it: the GPU batch replaces the process pool


def _n_steps(env, sim_time):
    """Env-steps taken by `while env.time < start + sim_time` (sim_engine.py:33): env.time is start_time plus the
    patient's clock, which starts at T1DPatient(t0=...) (env.py:44-46, t1dpatient.py:27), so an episode that starts
    late takes ceil((minutes - t0) / sample_time) steps."""
    minutes = sim_time.total_seconds() / 60.0 - float(getattr(env.patient, 't0', 0))
    return max(0, int(np.ceil(minutes / float(env.sample_time))))


def _meal_table(env, n_minutes):
    """Replay scenario.get_action once per episode minute, as env.mini_step would (env.py:50)."""
    sc = env.scenario
    # this package's own scenario classes build the same table (and leave the same generator state) day by day
    from .scenario import CustomScenario
    from .scenario_gen import RandomScenario
    for cls in (RandomScenario, CustomScenario):
        if isinstance(sc, cls) and type(sc).get_action is cls.get_action:
            fast = sc.meal_table(env.patient.t0, n_minutes)
            if fast is not None:
                return fast
    start = env.scenario.start_time
    mm, gg = [], []
    for m in range(int(n_minutes)):
        g = env.scenario.get_action(start + timedelta(minutes=env.patient.t0 + m)).meal
        if g != 0:
            mm.append(m + env.patient.t0)
            gg.append(float(g))
    return np.array(mm, dtype=np.int32), np.array(gg, dtype=np.float64)


class SimObj(object):
    def __init__(self, env, controller, sim_time, animate=True, path=None):
        self.env = env
        self.controller = controller
        self.sim_time = sim_time
        self.animate = animate
        self._ctrller_kwargs = None
        self.path = path

    def _device_plan(self):
        """(kind, ctrl) if this simulation can run as a fused episode, else None."""
        spec = getattr(self.controller, 'device_spec', None)
        if not getattr(self.env, 'fused', False) or spec is None:
            return None
        return spec(self.env.patient.name)

    def simulate(self):
        tic = time.time()
        if self._device_plan() is not None:
            _run_fused([self])
        else:
            self._simulate_lockstep()
        logger.info('Simulation took {} seconds.'.format(time.time() - tic))

    def _simulate_lockstep(self):
        """The reference's loop: controller.policy on the host, one env-step launch per step."""
        self.controller.reset()
        obs, reward, done, info = self.env.reset()
        while self.env.time < self.env.scenario.start_time + self.sim_time:
            if self.animate:
                self.env.render()
            action = self.controller.policy(obs, reward, done, **info)
            obs, reward, done, info = self.env.step(action)

    def results(self):
        return self.env.show_history()

    def save_results(self):
        df = self.results()
        if not os.path.isdir(self.path):
            os.makedirs(self.path)
        df.to_csv(os.path.join(self.path, str(self.env.patient.name) + '.csv'))

    def reset(self):
        self.env.reset()
        self.controller.reset()


def _run_fused(sims):
    """All `sims` (same hardware, controller kind and step count) as one population, one episode launch."""
    envs = [s.env for s in sims]
    E = len(sims)
    first = envs[0]
    n_steps = _n_steps(first, sims[0].sim_time)
    st = int(first.sample_time)
    kinds, ctrl_rows = [], np.zeros((4, E))
    for j, s in enumerate(sims):
        s.controller.reset()
        kind, ctrl = s._device_plan()
        kinds.append(kind)
        ctrl_rows[:len(ctrl), j] = ctrl
        s.env.reset()                                # patient / sensor / pump / scenario resets + history row 0
    assert len(set(kinds)) == 1
    tables = [_meal_table(e, n_steps * st) for e in envs]
    # (the replay above leaves every scenario object in the state the reference's loop would: same calls)
    mm, mg = streams.pad_tables(tables)
    noise = np.stack([e.sensor._noise_generator.peek(n_steps + 2) for e in envs], axis=1)
    x0 = np.stack([np.asarray(e.patient.init_state, dtype=np.float64) for e in envs], axis=1)
    params = np.stack([_hot(e.patient) for e in envs], axis=1)
    p0 = first.patient
    eng = BatchEngine(E, device=p0._engine.device.index, carry_h=bool(p0._engine.flags & 1),
                      fma=bool(p0._engine.flags & 2), sensor=dict(first.sensor._params), pump=dict(first.pump._params))
    try:
        eng.set_t0(p0.t0)
        eng.reset(np.zeros(E, dtype=np.int32), noise, mm, mg, x0=x0, params=params, row0=False)
        out = eng.run(n_steps, controller=kinds[0], ctrl_rows=ctrl_rows, trace=True)
        trace = out['trace'].permute(2, 0, 1).contiguous().cpu().numpy()        # [E, n_steps, NTRACE]
        d, i = eng.episode_state()
        d_host = d.cpu().numpy()
        status = eng.status().cpu().numpy()
        for j, e in enumerate(envs):
            # hand the episode back to the patient's own handle so that the env can be stepped further
            pe = e.patient._engine
            if n_steps + 3 > e._n_noise:
                e._n_noise = n_steps + 2 + 1024
                pe.rebind_streams(e._noise_tensor(e._n_noise))
            pe.set_episode_state(d[:, j:j + 1].contiguous(), i[:, j:j + 1].contiguous(), steps_done=n_steps)
            e._sync_plant()
            e._append_trace(trace[j])
            if status[j] < 0:
                raise RuntimeError('ODE solver failed for %s (dopri5 idid = %d)' % (e.patient.name, status[j]))
            if kinds[0] == 'pid':                    # SG_D_PID_INTEG, SG_D_PID_PREV
                sims[j].controller.integrated_state = float(d_host[22, j])
                sims[j].controller.prev_state = float(d_host[23, j])
    finally:
        eng.close()


def _hot(patient):
    from ..patient.t1dpatient import _hot_row
    return _hot_row(patient._params)


def _group_key(s):
    plan = s._device_plan()
    if plan is None:
        return None
    e = s.env
    return (plan[0], _n_steps(e, s.sim_time), int(e.sample_time), tuple(sorted(dict(e.sensor._params).items(), key=str)),
            tuple(sorted(dict(e.pump._params).items(), key=str)), e.patient.t0, e.patient._engine.flags,
            e.patient._engine.device.index)


def sim(sim_object):
    print("Process ID: {}".format(os.getpid()))
    print('Simulation starts ...')
    sim_object.simulate()
    sim_object.save_results()
    print('Simulation Completed!')
    return sim_object.results()


def batch_sim(sim_instances, parallel=False):
    """Run every SimObj, save its CSV, return the list of result DataFrames (in order).  `parallel` is
    accepted for compatibility: fusable simulations always run together on the GPU."""
    tic = time.time()
    groups, loose = {}, []
    for idx, s in enumerate(sim_instances):
        k = _group_key(s)
        (groups.setdefault(k, []) if k is not None else loose).append(idx)
    print('Simulation starts ...')
    for k, members in groups.items():
        _run_fused([sim_instances[i] for i in members])
    for i in loose:
        sim_instances[i]._simulate_lockstep()
    results = []
    for s in sim_instances:
        if s.path is not None:
            s.save_results()
        results.append(s.results())
    print('Simulation Completed!')
    print('Simulation took {} sec.'.format(time.time() - tic))
    return results
