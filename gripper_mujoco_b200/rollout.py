"""Vectorised DQN rollout and replay over the batched MjClass (the immediate caller of the
hot path). Mirrors `Agent_DQN` / `ReplayMemory` (rl/agents/DQN.py:11-393) and `Trainer.run_episode`
(rl/Trainer.py:642-708) with a leading environment axis; nothing in the loop leaves the device:

* `ReplayRing`   -- `ReplayMemory`: the reference keeps a `deque(maxlen=capacity)` of per-transition tensors and
  draws with `random.Random.sample` (rl/agents/DQN.py:11-46). Here the transitions live in preallocated device
  tensors, a push writes a whole batch at the ring cursor (same FIFO eviction as the deque) and a draw is one
  index gather. Sampling is uniform without replacement like the reference's; the generator differs (torch).
* `VectorDQN`    -- `Agent_DQN`: same `Parameters` (names and defaults), same epsilon schedule per env
  (`eps_end + (eps_start - eps_end) exp(-decay_num / eps_decay)`, decay_num = that env's episode number), same
  loss / optimiser / gradient clamp / soft or hard target update. The reference optimises once per pushed
  transition; with n envs a push is n transitions, so `updates_per_push` sets how many batches are drawn per
  vector step (`None` keeps the reference's 1 : 1 ratio).
  Quirk kept: the reference never passes `next_state=None`, so its non-final mask is all true and the target
  bootstraps through terminal states (rl/agents/DQN.py:262-293 with rl/Trainer.py:664-684);
  `mask_terminal=True` switches to the textbook target.
* `VectorRollout` -- `Trainer.run_episode` for all envs at once: act, `fused_step`, push, optimise, masked
  `reset` + `spawn_into_scene` of finished episodes (`MjEnv.step` / `MjEnv.reset`, rl/env/MjEnv.py:2170-2263,
  truncation at `max_episode_steps`, :616-637), per-episode return / length log.

The policy network is an `nn.Module` with `n_output` like the reference's (rl/networks.py:7-40); its GEMMs are
library calls (cuBLAS through torch), as they are in the reference.
"""
from __future__ import annotations

import math
from copy import deepcopy
from dataclasses import asdict, dataclass

import torch
import torch.nn as nn


class ReplayRing:
    """Device-resident FIFO ring of transitions (state, action, next_state, reward, terminal)."""

    def __init__(self, capacity: int, n_obs: int, device, rngseed: int | None = None):
        self.capacity, self.n_obs, self.device = int(capacity), int(n_obs), torch.device(device)
        self.state = torch.zeros((capacity, n_obs), dtype=torch.float32, device=self.device)
        self.next_state = torch.zeros((capacity, n_obs), dtype=torch.float32, device=self.device)
        self.action = torch.zeros(capacity, dtype=torch.int64, device=self.device)
        self.reward = torch.zeros(capacity, dtype=torch.float32, device=self.device)
        self.terminal = torch.zeros(capacity, dtype=torch.bool, device=self.device)
        self.cursor = 0          # next write position
        self.size = 0
        self.pushed = 0          # transitions ever pushed
        self.seed(rngseed)

    def seed(self, rngseed: int | None = None) -> None:
        self.rngseed = rngseed
        self.rng = torch.Generator(device=self.device)
        if rngseed is not None:
            self.rng.manual_seed(int(rngseed))

    def __len__(self) -> int:
        return self.size

    def push(self, state, action, next_state, reward, terminal) -> None:
        """Append a batch of n transitions in order; the oldest entries are overwritten once full, exactly as
        n successive `deque.append` calls would evict them."""
        n = state.shape[0]
        if n > self.capacity:              # only the last `capacity` survive n appends
            sl = slice(n - self.capacity, n)
            state, action, next_state, reward, terminal = state[sl], action[sl], next_state[sl], reward[sl], terminal[sl]
            self.pushed += n - self.capacity
            n = self.capacity
        idx = (torch.arange(n, device=self.device) + self.cursor) % self.capacity
        self.state.index_copy_(0, idx, state.to(self.device, torch.float32))
        self.next_state.index_copy_(0, idx, next_state.to(self.device, torch.float32))
        self.action.index_copy_(0, idx, action.to(self.device, torch.int64).reshape(-1))
        self.reward.index_copy_(0, idx, reward.to(self.device, torch.float32).reshape(-1))
        self.terminal.index_copy_(0, idx, terminal.to(self.device, torch.bool).reshape(-1))
        self.cursor = (self.cursor + n) % self.capacity
        self.size = min(self.capacity, self.size + n)
        self.pushed += n

    def ordered(self):
        """Contents oldest -> newest (what `list(deque)` holds in the reference)."""
        if self.size < self.capacity:
            idx = torch.arange(self.size, device=self.device)
        else:
            idx = (torch.arange(self.capacity, device=self.device) + self.cursor) % self.capacity
        return self.state[idx], self.action[idx], self.next_state[idx], self.reward[idx], self.terminal[idx]

    def sample_indices(self, batch_size: int) -> torch.Tensor:
        if batch_size > self.size:
            raise ValueError("Sample larger than population")          # random.sample raises the same
        # uniform without replacement: top-k of iid keys (no host sync, unlike randperm on `size`)
        keys = torch.rand(self.size, device=self.device, generator=self.rng)
        return torch.topk(keys, batch_size, sorted=False).indices

    def batch(self, batch_size: int):
        idx = self.sample_indices(batch_size)
        return self.state[idx], self.action[idx], self.next_state[idx], self.reward[idx], self.terminal[idx]

    def state_dict(self) -> dict:
        s, a, ns, r, t = self.ordered()
        return {"state": s.cpu(), "action": a.cpu(), "next_state": ns.cpu(), "reward": r.cpu(), "terminal": t.cpu(),
                "capacity": self.capacity, "pushed": self.pushed}

    def load_state_dict(self, d: dict) -> None:
        self.cursor = self.size = 0
        pushed = d.get("pushed", len(d["action"]))
        if len(d["action"]):
            self.push(d["state"], d["action"], d["next_state"], d["reward"], d["terminal"])
        self.pushed = pushed


class VectorDQN:
    """`Agent_DQN` (rl/agents/DQN.py:64-393) over n environments at once."""

    name = "Agent_DQN"

    @dataclass
    class Parameters:                      # rl/agents/DQN.py:68-98, same names and defaults
        learning_rate: float = 1e-4
        gamma: float = 0.99
        batch_size: int = 128
        eps_start: float = 0.9
        eps_end: float = 0.05
        eps_decay: int = 1000
        min_memory_replay: int = 250
        memory_replay: int = 10_000
        optimiser: str = "adamW"
        adam_beta1: float = 0.9
        adam_beta2: float = 0.999
        target_update: int = 250
        soft_target_update: bool = True
        soft_target_tau: float = 0.05
        use_double_dqn: bool = False
        grad_clamp_value: float | None = 100
        loss_criterion: str = "smoothL1Loss"

        def update(self, newdict):
            for key, value in newdict.items():
                if hasattr(self, key):
                    setattr(self, key, value)
                else:
                    raise RuntimeError(f"incorrect key: {key}")

    def __init__(self, device="cpu", rngseed=None, mode="train", updates_per_push: int | None = None,
                 mask_terminal: bool = False):
        self.params = VectorDQN.Parameters()
        self.device = torch.device(device)
        self.rngseed, self.mode = rngseed, mode
        self.updates_per_push, self.mask_terminal = updates_per_push, mask_terminal
        self.n_updates = 0

    def init(self, network, n_obs: int | None = None):
        p = self.params
        if network != "loaded":
            self.policy_net = network
            self.target_net = deepcopy(network)
            self.n_actions = network.n_output
        self.policy_net.to(self.device); self.target_net.to(self.device)
        if n_obs is None:
            n_obs = next(self.policy_net.parameters()).shape[1]
        self.memory = ReplayRing(p.memory_replay, n_obs, self.device)
        opt = p.optimiser.lower()
        if opt == "rmsprop":
            self.optimiser = torch.optim.RMSprop(self.policy_net.parameters(), lr=p.learning_rate)
        elif opt == "adam":
            self.optimiser = torch.optim.Adam(self.policy_net.parameters(), lr=p.learning_rate,
                                              betas=(p.adam_beta1, p.adam_beta2))
        elif opt == "adamw":
            self.optimiser = torch.optim.AdamW(self.policy_net.parameters(), lr=p.learning_rate,
                                               betas=(p.adam_beta1, p.adam_beta2), amsgrad=True)
        else:
            raise RuntimeError(f"Invalid optimiser choice '{p.optimiser}', options are 'adam' or 'rmsprop'")
        crit = p.loss_criterion.lower()
        if crit == "smoothl1loss":
            self.loss_criterion = nn.SmoothL1Loss()
        elif crit == "mseloss":
            self.loss_criterion = nn.MSELoss()
        elif crit == "huber":
            self.loss_criterion = nn.HuberLoss()
        else:
            raise RuntimeError(f"Invalid loss criterion '{p.loss_criterion}'")
        self.seed()
        if self.mode == "train":
            self.training_mode()
        elif self.mode == "test":
            self.testing_mode()
        else:
            raise RuntimeError(f"Agent_DQN given mode={self.mode} but valid options are 'test' and 'train'")

    def seed(self, rngseed=None):
        if rngseed is None:
            rngseed = self.rngseed if self.rngseed is not None else int(torch.randint(0, 2_147_483_647, (1,)).item())
        self.rngseed = rngseed
        self.rng = torch.Generator(device=self.device); self.rng.manual_seed(int(rngseed))
        self.memory.seed(rngseed)

    def eps_threshold(self, decay_num: torch.Tensor) -> torch.Tensor:
        p = self.params
        return p.eps_end + (p.eps_start - p.eps_end) * torch.exp(-1.0 * decay_num.to(torch.float32) / float(p.eps_decay))

    def select_action(self, state: torch.Tensor, decay_num, test: bool = False) -> torch.Tensor:
        """state [n, n_obs] -> int64 actions [n]; decay_num is a scalar or an [n] tensor of episode numbers
        (rl/agents/DQN.py:186-213)."""
        n = state.shape[0]
        with torch.no_grad():
            greedy = self.policy_net(state.to(self.device)).max(1)[1]
        if test:
            return greedy
        if not torch.is_tensor(decay_num):
            decay_num = torch.full((n,), float(decay_num), device=self.device)
        rand = torch.rand(n, device=self.device, generator=self.rng)
        explore = rand < self.eps_threshold(decay_num.to(self.device))
        random_actions = torch.randint(0, self.n_actions, (n,), device=self.device, generator=self.rng)
        return torch.where(explore, random_actions, greedy)

    def td_loss(self, state, action, next_state, reward, terminal) -> torch.Tensor:
        """The loss of one `optimise_model` batch (rl/agents/DQN.py:262-303)."""
        p = self.params
        q = self.policy_net(state).gather(1, action.reshape(-1, 1))
        with torch.no_grad():
            if p.use_double_dqn:
                nxt_a = self.policy_net(next_state).max(1)[1].unsqueeze(1)
                nxt = self.target_net(next_state).gather(1, nxt_a).squeeze(1)
            else:
                nxt = self.target_net(next_state).max(1)[0]
            if self.mask_terminal:
                nxt = torch.where(terminal, torch.zeros_like(nxt), nxt)
        expected = nxt * p.gamma + reward
        return self.loss_criterion(q, expected.unsqueeze(1))

    def optimise_on(self, state, action, next_state, reward, terminal) -> torch.Tensor:
        loss = self.td_loss(state, action, next_state, reward, terminal)
        self.optimiser.zero_grad()
        loss.backward()
        if self.params.grad_clamp_value is not None:
            torch.nn.utils.clip_grad_value_(self.policy_net.parameters(), self.params.grad_clamp_value)
        self.optimiser.step()
        self.n_updates += 1
        return loss.detach()

    def optimise_model(self):
        p = self.params
        if len(self.memory) < p.batch_size or len(self.memory) < p.min_memory_replay:
            return None
        return self.optimise_on(*self.memory.batch(p.batch_size))

    def soft_update(self) -> None:
        tau = self.params.soft_target_tau
        with torch.no_grad():
            for tp, pp in zip(self.target_net.state_dict().values(), self.policy_net.state_dict().values()):
                if tp.is_floating_point():
                    tp.mul_(1 - tau).add_(pp, alpha=tau)
                else:
                    tp.copy_(pp)

    def update_step(self, state, action, next_state, reward, terminal, truncated=None):
        """One vector step: push n transitions, then `updates_per_push` (default n) optimise + soft-target rounds
        (rl/agents/DQN.py:329-360 runs one round per transition)."""
        if self.mode != "train":
            return
        n = state.shape[0]
        self.memory.push(state, action, next_state, reward, terminal)
        rounds = n if self.updates_per_push is None else self.updates_per_push
        for _ in range(rounds):
            self.optimise_model()
            if self.params.soft_target_update:
                self.soft_update()

    def update_episode(self, i_episode: int, finished: bool = False, n_finished: int = 1):
        """rl/agents/DQN.py:362-377; with n envs, `n_finished` episodes end at once: the hard target update fires
        if any of the episode numbers (i_episode - n_finished, i_episode] is a multiple of target_update."""
        if finished:
            self.target_net.load_state_dict(self.policy_net.state_dict())
        if self.mode != "train" or self.params.soft_target_update:
            return
        tu = self.params.target_update
        if i_episode // tu > (i_episode - n_finished) // tu:
            self.target_net.load_state_dict(self.policy_net.state_dict())

    def training_mode(self):
        self.mode = "train"; self.policy_net.train(); self.target_net.train()

    def testing_mode(self):
        self.mode = "test"; self.policy_net.eval(); self.target_net.eval()

    def get_params_dict(self):
        d = asdict(self.params); d.update({"network_name": getattr(self.policy_net, "name", type(self.policy_net).__name__)})
        return d

    def get_save_state(self):
        return {"name": self.name, "parameters": self.params, "memory": self.memory.state_dict(),
                "network": self.policy_net, "optimiser_state_dict": self.optimiser.state_dict()}

    def load_save_state(self, saved, device=None):
        if self.name != saved["name"]:
            raise RuntimeError(f"{self.name}.load_save_state() failed as given save state has agent name: {saved['name']}")
        self.params = saved["parameters"]
        if device is not None:
            self.device = torch.device(device)
        self.init(saved["network"])
        self.memory.load_state_dict(saved["memory"])
        self.optimiser.load_state_dict(saved["optimiser_state_dict"])


class VariableNetwork(nn.Module):
    """The reference's MLP Q-network (rl/networks.py:7-40): Linear+ReLU stack with a softmax head."""

    name = "VariableNetwork_"

    def __init__(self, layers, device="cpu"):
        super().__init__()
        self.device = device
        self.n_input, self.n_output = layers[0], layers[-1]
        self.linear = nn.ModuleList([nn.Linear(layers[i], layers[i + 1]) for i in range(len(layers) - 1)])
        self.name = "VariableNetwork_" + "x".join(str(v) for v in layers[1:-1])
        self.activation = nn.ReLU()
        self.softmax = nn.Softmax(dim=1)

    def forward(self, x):
        x = x.to(self.device)
        for lin in self.linear[:-1]:
            x = self.activation(lin(x))
        return self.softmax(self.linear[-1](x))


class VectorRollout:
    """`Trainer.run_episode` / `Trainer.train` over every env of a batched `MjClass` at once."""

    def __init__(self, mj, agent: VectorDQN, max_episode_steps: int = 250, spawn_xy_range: float = 0.01,
                 spawn_rot_range: float = math.pi / 2):
        self.mj, self.agent, self.max_episode_steps = mj, agent, int(max_episode_steps)
        self.spawn_xy_range, self.spawn_rot_range = spawn_xy_range, spawn_rot_range
        dev, n = mj.device, mj.n_envs
        self.ep_steps = torch.zeros(n, dtype=torch.int32, device=dev)
        self.ep_return = torch.zeros(n, dtype=torch.float32, device=dev)
        self.ep_index = torch.ones(n, dtype=torch.int64, device=dev)       # i_episode of each env (starts at 1)
        self.episodes_done = 0
        self.steps_done = 0
        self.returns_sum = torch.zeros((), dtype=torch.float64, device=dev)
        self.lengths_sum = torch.zeros((), dtype=torch.float64, device=dev)
        self.obs = None

    def reset_all(self):
        self.mj.reset()
        self.mj.batch.spawn_with_retry(self.spawn_xy_range, self.spawn_rot_range)
        self.obs = self.mj.get_observation().clone()
        self.ep_steps.zero_(); self.ep_return.zero_()
        return self.obs

    def step(self, test: bool = False):
        """One action-step of every env; returns (reward, terminated, truncated) device tensors."""
        if self.obs is None:
            self.reset_all()
        gb = self.mj.batch
        action = self.agent.select_action(self.obs, self.ep_index, test=test)
        obs, reward, done, _ = self.mj.fused_step(action.to(torch.int32))
        self.ep_steps.add_(1)
        terminated = done != 0
        truncated = (self.ep_steps >= self.max_episode_steps) & ~terminated     # MjEnv._is_done checks the cap first
        terminated = terminated & (self.ep_steps < self.max_episode_steps)
        finished = terminated | truncated
        new_obs = obs.clone()
        if not test:
            self.agent.update_step(self.obs, action, new_obs, reward, finished, truncated)
        self.ep_return.add_(reward)
        # episode bookkeeping + masked MjEnv.reset (reset -> spawn) for the finished envs
        self.returns_sum += (self.ep_return * finished).sum(dtype=torch.float64)
        self.lengths_sum += (self.ep_steps * finished).sum(dtype=torch.float64)
        m8 = finished.to(torch.uint8)
        gb.reset(m8); gb.spawn_with_retry(self.spawn_xy_range, self.spawn_rot_range, m8)
        fresh = self.mj.get_observation()
        self.obs = torch.where(finished.unsqueeze(1), fresh, new_obs)
        keep = (~finished)
        self.ep_steps.mul_(keep.to(torch.int32)); self.ep_return.mul_(keep.to(torch.float32))
        self.ep_index.add_(finished.to(torch.int64))
        self.steps_done += 1
        return reward, terminated, truncated

    def run(self, n_steps: int, test: bool = False, sync_every: int = 0) -> dict:
        """n_steps vector steps; episode counters are read back once at the end (or every `sync_every` steps,
        which is also when a hard target update can fire)."""
        done_before = int((self.ep_index - 1).sum().item())
        for k in range(n_steps):
            self.step(test=test)
            if sync_every and (k + 1) % sync_every == 0:
                self._sync_episodes()
        self._sync_episodes()
        total = self.episodes_done
        new = total - done_before
        return {"vector_steps": n_steps, "transitions": n_steps * self.mj.n_envs, "episodes": new,
                "mean_return": float(self.returns_sum.item()) / max(1, total),
                "mean_length": float(self.lengths_sum.item()) / max(1, total),
                "updates": self.agent.n_updates, "memory": len(self.agent.memory)}

    def _sync_episodes(self):
        total = int((self.ep_index - 1).sum().item())
        n_new = total - self.episodes_done
        if n_new > 0:
            self.agent.update_episode(total, n_finished=n_new)
        self.episodes_done = total
