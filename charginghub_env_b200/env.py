"""The reference's gym surface on top of the CUDA engine.

EvcsspManagerEnv_v6  -- drop-in for evcssp_env_cpp.envs.evcssp_manager.EvcsspManagerEnv_v6
                        (same constructor kwargs, reset() -> obs, step(a) -> (obs, reward, done, {}),
                        observation/action spaces, diagnostic attributes); one hub, numpy in/out.
BatchedHubEnv        -- the same environment for H hubs at once: torch tensors in and out (DLPack
                        views of engine memory), fused auto-reset, optional CUDA-graph rollouts.
Both only marshal data; every hub is advanced by the sm_100a kernels behind include/chub.h.
"""
import time as _time

import numpy as np

from .engine import HubEngine, load_tables
from .exogenous import HubExogenous, synthetic_year

try:  # gym 0.18.3 is what the reference targets; neither it nor gymnasium is required
    import gym as _gym
    _EnvBase, _spaces = _gym.Env, _gym.spaces
except Exception:  # pragma: no cover - depends on the image
    try:
        import gymnasium as _gym
        _EnvBase, _spaces = _gym.Env, _gym.spaces
    except Exception:
        _gym, _EnvBase, _spaces = None, object, None


class Box(object):
    """Minimal stand-in for gym.spaces.Box when no gym is installed."""

    def __init__(self, low, high, shape=None, dtype=np.float32):
        self.dtype = np.dtype(dtype)
        self.shape = tuple(shape) if shape is not None else np.shape(low)
        self.low = np.broadcast_to(np.asarray(low, dtype=dtype), self.shape).copy()
        self.high = np.broadcast_to(np.asarray(high, dtype=dtype), self.shape).copy()

    def sample(self):
        return np.random.uniform(self.low, self.high).astype(self.dtype)

    def contains(self, x):
        x = np.asarray(x)
        return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))


def _box(low, high, shape=None, dtype=np.float32):
    if _spaces is not None:
        return _spaces.Box(low=low, high=high, shape=shape, dtype=dtype) if shape is not None else \
            _spaces.Box(low=np.asarray(low, dtype), high=np.asarray(high, dtype), dtype=dtype)
    return Box(low, high, shape, dtype)


def observation_bounds(station_list, price_table):
    """low/high of the observation (evcssp_manager.py:48-104)."""
    obs_price = (np.asarray(price_table) - np.mean(price_table)) / np.std(price_table)
    lo, hi = [-1.0, float(min(obs_price))], [1.0, float(max(obs_price))]
    n_st = 1 if (station_list[0] == 0 or station_list[1] == 0) else 2
    for _ in range(n_st):
        lo += [-1.0, -1.0, -1.0, 0]
        hi += [1.0, 1.0, 1.0, 2]
    lo += [0, 0, 0]
    hi += [1, 1, 1]
    return np.array(lo, np.float32), np.array(hi, np.float32)


class EvcsspManagerEnv_v6(_EnvBase):
    metadata = {"render.modes": ["human", "rgb_array"], "video.frames_per_second": 30}

    def __init__(self, station_list, station_type_list, constant_charging=False, hydro_prod_rate=None,
                 hydro_store_vlt=None, seed_rand=True, init_soc=0.5, fc_max_power=None, fcev_permeate=0.01,
                 use_lagrange=False, renew_fluctuate=0, price_fluctuate=0, hydro_loss=0, device=0, el_table=None):
        assert len(station_list) == len(station_type_list) == 2
        self._kw = dict(station_list=list(station_list), station_type_list=list(station_type_list),
                        constant_charging=constant_charging, hydro_prod_rate=hydro_prod_rate,
                        hydro_store_vlt=hydro_store_vlt, init_soc=init_soc, fc_max_power=fc_max_power,
                        fcev_permeate=fcev_permeate, hydro_loss=hydro_loss, el_table=el_table)
        self.tables = load_tables()
        # Change_Use_Seed(seed_rand) (CHS.hpp:27-44): True seeds the stream from the clock, False is repeatable
        base_seed = int(_time.time_ns()) if seed_rand else 0
        self.engine = HubEngine(1, [self._kw], device=device, diagnostics=True, tables=self.tables, seed=base_seed)
        self.pile_number = list(station_list)
        self._n_act = sum(station_list) + 2
        self._n_obs = 9 if (station_list[0] == 0 or station_list[1] == 0) else 13
        self.exo = HubExogenous(self.tables, renew_fluctuate, price_fluctuate)
        self.engine.set_traces(np.zeros(2), np.zeros(2), np.zeros(2))  # 2-slot ring poked every step
        self.price_mean, self.price_std = self.tables["price_mean"], self.tables["price_std"]
        self.real_time_max = 96
        self.low_state, self.high_state = observation_bounds(station_list, self.tables["price"])
        self.action_space = _box(-1.0, 1.0, shape=(self._n_act,), dtype=np.float32)
        self.observation_space = _box(self.low_state, self.high_state, dtype=np.float32)
        self.hy_init_soc = init_soc
        self.np_random = None
        self.seed()
        self._k = 0
        self._time = 0
        self.reset()
        self.real_state, self.state, self.np_random = [], None, None  # evcssp_manager.py:121-123
        self.acumulate_reward = 0
        self.cumulated_income = 0
        self.cumulated_draw_ele = 0
        self.use_lagrangian, self.lagrangian_factor, self.penalty = False, None, 0

    def seed(self, seed=None):
        self.np_random = np.random.RandomState(seed)
        return [seed]

    # the reference's state vector before normalisation (evcssp_manager.py:364-372)
    def _real_state(self, pv, wd, price, rep=None):
        rep = rep or self.engine.hub_report(0)   # every scalar of the hub in one device round trip
        rs = [float(self._time), price]
        for s in range(2):
            if self.pile_number[s] > 0:
                rs += [rep["st_min_power"][s], rep["st_charge_power"][s], rep["st_max_power"][s], rep["st_line"][s]]
        rs += [rep["store_soc"], pv, wd]
        return np.array(rs)

    def reset(self):
        self.cumulated_income = 0
        self.cumulated_draw_ele = 0
        self.exo.new_day()
        pv, wd, price = self.exo.state_values(0, float(self.tables["price"][95]))
        self._k = 0
        self.engine.poke_trace(0, pv, wd, price)
        obs = self.engine.reset_host()
        self.exo.reset_price_count()
        self._time = 0
        self.re_pv_power, self.re_wd_power = pv, wd
        self.real_state = self._real_state(pv, wd, price)
        self.state = np.array(obs[0, : self._n_obs], dtype=np.float64)
        return np.array(self.state)

    def step(self, action):
        if action is None:  # evcssp_manager.py:146-147
            action = np.append(np.ones(sum(self.pile_number)), [0, 0], None)
        assert len(action) == self._n_act
        t_next = (self._time + 1) % 96
        # exogenous values of the state being entered: price list's last entry is table[t] (Aggregator_Simple.py:147)
        pv, wd, price = self.exo.state_values(t_next, float(self.tables["price"][self._time]))
        self.engine.poke_trace((self._k + 1) % 2, pv, wd, price)
        obs, rew, done = self.engine.step_host(np.asarray(action, np.float32).reshape(1, -1))
        self._k += 1
        self._time = t_next
        rep = self.engine.hub_report(0)
        reward = rep["reward"]
        self.income = rep["income"]
        self.fc_power = rep["fc_power"]
        self.re_hy_for_fc = rep["hy_to_use"]
        self.re_hydrogen_power = rep["h2_grid"]
        self.re_hydrogen_power_init = rep["h2_power"]
        self.re_ev_power_list = [rep["ev_grid0"], rep["ev_grid1"]]
        self.re_ev_power_sum = self.real_charging_power = sum(self.re_ev_power_list)
        self.hy_act = rep["hy_act"]
        self.re_hy_gen = 900.0 * rep["hy_flow"]
        self.cumulated_income += self.income
        self.cumulated_draw_ele += self.re_ev_power_list[0] + self.re_ev_power_list[1] + self.re_hydrogen_power
        self.acumulate_reward += reward
        store_soc = rep["store_soc"]
        self.deviation = abs(store_soc - self.hy_init_soc)
        if bool(done[0]):  # evcssp_manager.py:275-290: computed and kept, the reward deduction itself is commented out
            cap_g = 0.089 * 200 * (5000.0 if self._kw["hydro_store_vlt"] is None else self._kw["hydro_store_vlt"]) * 1000
            dev = (store_soc - self.hy_init_soc) if self.use_lagrangian else abs(store_soc - self.hy_init_soc)
            self.test_penalty = abs(dev * cap_g / 1000 / 0.2)
        self.re_pv_power, self.re_wd_power = pv, wd
        self.real_state = self._real_state(pv, wd, price, rep)
        self.state = np.array(obs[0, : self._n_obs], dtype=np.float64)
        return self.state, reward, bool(done[0]), {}

    def render(self, mode="human"):
        pass

    def close(self):
        if getattr(self, "engine", None) is not None:
            self.engine.close()

    def show_situation(self):
        cols, _ = self.engine.situation()
        for name in self.engine.SITUATION_COLUMNS:
            print(name, cols[name][0].cpu().numpy())


class BatchedHubEnv(object):
    """H hubs, torch in / torch out.  step() returns views of engine memory (zero copy through DLPack).
    traces: (pv, wd, price) arrays, None for the shared synthetic year, or "device" for the on-device generator."""

    def __init__(self, n_hubs, env_kwargs=None, classes=None, hub_class=None, device=0, hub_id_base=0,
                 auto_reset=True, seed=0, traces=None, action_stride=None, diagnostics=False):
        if classes is None:
            classes = [dict(env_kwargs or {})]
        fluct = [(c.get("renew_fluctuate", 0) or 0, c.get("price_fluctuate", 0) or 0) for c in classes]
        classes = [{k: v for k, v in c.items() if k not in ("seed_rand", "use_lagrange", "renew_fluctuate",
                                                              "price_fluctuate")} for c in classes]
        self.tables = load_tables()
        self.engine = HubEngine(n_hubs, classes, hub_class=hub_class, device=device, hub_id_base=hub_id_base,
                                auto_reset=auto_reset, diagnostics=diagnostics, tables=self.tables, seed=seed,
                                action_stride=action_stride)
        if isinstance(traces, str) and traces == "device":
            # the reference's own exogenous model, generated per hub inside the step kernel (random PV / wind days at
            # every reset, OU noise on both and on the price; renew_fluctuate / price_fluctuate of the first class)
            self.engine.set_exogenous_generator(self.tables, fluct[0][0], fluct[0][1])
        else:
            pv, wd, price = traces if traces is not None else synthetic_year(self.tables)
            self.engine.set_traces(pv, wd, price, per_hub=np.ndim(pv) == 2)
        self.n_hubs, self.action_dim = n_hubs, self.engine.action_stride
        lo, hi = observation_bounds([1, 1], self.tables["price"])
        self.observation_space = _box(lo, hi, dtype=np.float32)
        self.action_space = _box(-1.0, 1.0, shape=(self.action_dim,), dtype=np.float32)

    def reset(self, mask=None):
        return self.engine.reset(mask=mask, step_index=0 if mask is None else -1)

    def step(self, actions):
        obs, reward, done = self.engine.step(actions)
        return obs, reward, done, {}

    def step_loads(self, loads):
        """Upper-level control: one power target (kW) per station, rows [load0, load1, a[-2], a[-1]]."""
        obs, reward, done = self.engine.step_loads(loads)
        return obs, reward, done, {}

    def capture_rollout(self, actions, n_steps):
        self.engine.capture_rollout(actions, n_steps)

    def rollout(self):
        self.engine.rollout()
        return self.engine.tensor("obs"), self.engine.tensor("reward"), self.engine.tensor("done")

    def episode_stats(self):
        """[H, 4] of the last finished episode: return, income, grid energy drawn, incomplete departures."""
        return self.engine.tensor("last_stats")

    def close(self):
        self.engine.close()


class VectorHubEnv(object):
    """gymnasium.vector.VectorEnv-shaped adapter over BatchedHubEnv (same-step autoreset): reset() -> (obs, info),
    step(a) -> (obs, reward, terminated, truncated, info), everything CUDA tensors.

    max_episode_steps: the reference registers the env under a TimeLimit of 999 steps (evcssp_env_cpp/__init__.py:3-8),
        which never fires because an episode ends after 96; a smaller value truncates (masked reset) like TimeLimit would.
        The reward returned with the truncating step is that step's reward (a masked reset leaves reward / done alone),
        the cut episode's statistics land in info["episode_stats"], and traces stay aligned with each hub's own time of
        day (include/chub.h chub_set_traces).
    soc_penalty: the end-of-episode hydrogen-store penalty the reference computes but leaves commented out
        (evcssp_manager.py:275-297): None (as shipped), "abs" (reward -= test_penalty) or "lagrangian"
        (reward -= signed deviation * lagrangian_factor).  deviation = (SOC_end - init_soc) * capacity_kg / 0.2.
    """

    def __init__(self, n_hubs, env_kwargs=None, max_episode_steps=999, soc_penalty=None, lagrangian_factor=1.0, **kw):
        import torch
        if soc_penalty not in (None, "abs", "lagrangian"):
            raise ValueError("soc_penalty must be None, 'abs' or 'lagrangian'")
        self.env = BatchedHubEnv(n_hubs, env_kwargs=env_kwargs, auto_reset=True, **kw)
        self.engine = self.env.engine
        self.num_envs = n_hubs
        self.single_observation_space, self.single_action_space = self.env.observation_space, self.env.action_space
        self.observation_space = _box(np.tile(self.single_observation_space.low, (n_hubs, 1)),
                                      np.tile(self.single_observation_space.high, (n_hubs, 1)), dtype=np.float32)
        self.action_space = _box(-1.0, 1.0, shape=(n_hubs, self.env.action_dim), dtype=np.float32)
        self.max_episode_steps = int(max_episode_steps)
        self.soc_penalty, self.lagrangian_factor = soc_penalty, float(lagrangian_factor)
        classes = kw.get("classes") or [dict(env_kwargs or {})]
        dev = "cuda:%d" % self.engine.device
        hc = kw.get("hub_class")
        pick = torch.zeros(n_hubs, dtype=torch.long, device=dev) if hc is None else torch.as_tensor(
            np.asarray(hc), device=dev).long()
        vlt = [c.get("hydro_store_vlt") for c in classes]
        # capacity_mass = 0.089 * 200 * store_v * 1000 g (hydro_sys.py:94-98), / 1000 -> kg
        cap = torch.tensor([0.089 * 200 * (5000.0 if v is None else float(v)) for v in vlt], dtype=torch.float64, device=dev)
        init = torch.tensor([float(c.get("init_soc", 0.5)) for c in classes], dtype=torch.float64, device=dev)
        self._cap_kg, self._init_soc = cap[pick], init[pick]
        self._elapsed = torch.zeros(n_hubs, dtype=torch.int32, device=dev)

    def reset(self, seed=None, options=None):
        if seed is not None:
            self.engine.seed(int(seed))
        self._elapsed.zero_()
        return self.env.reset(), {}

    def step(self, actions):
        import torch
        obs, reward, done, _ = self.env.step(actions)
        terminated = done.bool()
        self._elapsed += 1
        info = {"final_soc": self.engine.tensor("final_soc"), "episode_stats": self.env.episode_stats()}
        if self.soc_penalty is not None:
            dev = (info["final_soc"].double() - self._init_soc) * self._cap_kg / 0.2
            pen = dev * self.lagrangian_factor if self.soc_penalty == "lagrangian" else dev.abs()
            pen = torch.where(terminated, pen, torch.zeros_like(pen)).float()
            reward = reward - pen
            info["soc_penalty"] = pen
        if self.max_episode_steps < 96:
            truncated = (self._elapsed >= self.max_episode_steps) & ~terminated
            obs = self.env.reset(mask=truncated.to(torch.uint8))
        else:
            truncated = torch.zeros_like(terminated)
        self._elapsed = torch.where(terminated | truncated, torch.zeros_like(self._elapsed), self._elapsed)
        return obs, reward, terminated, truncated, info

    def close(self):
        self.env.close()


def register():
    """Register `charging-hub-v6` with gym / gymnasium when one is installed (evcssp_env_cpp/__init__.py:3-8)."""
    if _gym is None:
        return False
    try:
        from gym.envs.registration import register as _reg
    except Exception:
        from gymnasium.envs.registration import register as _reg
    try:
        _reg(id="charging-hub-v6", entry_point="charginghub_env_b200.env:EvcsspManagerEnv_v6",
             max_episode_steps=999, reward_threshold=99)
    except Exception:
        return False
    return True
