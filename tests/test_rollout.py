"""Vectorised DQN rollout / replay (gripper_mujoco_b200/rollout.py) against golden vectors made by the
reference's own Agent_DQN / ReplayMemory (tests/golden/make_dqn_golden.py), plus the GPU rollout over the CUDA path."""
import os
from collections import deque

import numpy as np
import pytest
import torch

from gripper_mujoco_b200.rollout import ReplayRing, VariableNetwork, VectorDQN, VectorRollout

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "dqn_learner.npz"))


def golden_net():
    net = VariableNetwork([int(v) for v in G["layers"]], "cpu")
    net.load_state_dict({k[3:]: torch.tensor(G[k]) for k in G.files if k.startswith("w0/")})
    return net


@pytest.mark.parametrize("case", range(int(G["n_cases"])))
def test_learner_matches_reference_agent(case):
    """Three update_step rounds on a pinned batch: policy and target weights equal the reference's."""
    opt, crit, double, soft = G[f"case{case}/cfg"]
    ag = VectorDQN(device="cpu", rngseed=3, updates_per_push=1)
    n = len(G["action"])
    ag.params.update({"optimiser": str(opt), "loss_criterion": str(crit), "use_double_dqn": bool(int(double)),
                      "soft_target_update": bool(int(soft)), "batch_size": n, "min_memory_replay": n,
                      "memory_replay": 64, "learning_rate": 1e-2, "grad_clamp_value": 0.5, "gamma": 0.97})
    ag.init(golden_net())
    batch = tuple(torch.tensor(G[k]) for k in ("state", "action", "next_state", "reward", "terminal"))
    ag.memory.push(*batch)
    ag.memory.batch = lambda bs: batch
    for _ in range(3):
        ag.update_step(batch[0][:1], batch[1][:1], batch[2][:1], batch[3][:1], batch[4][:1])
    for k, v in ag.policy_net.state_dict().items():
        np.testing.assert_allclose(v.numpy(), G[f"case{case}/policy/{k}"], rtol=2e-5, atol=2e-6, err_msg=f"policy {k}")
    for k, v in ag.target_net.state_dict().items():
        np.testing.assert_allclose(v.numpy(), G[f"case{case}/target/{k}"], rtol=2e-5, atol=2e-6, err_msg=f"target {k}")


def test_epsilon_schedule_matches_reference():
    ag = VectorDQN(device="cpu", rngseed=1)
    eps = ag.eps_threshold(torch.tensor(G["eps_decay_num"]))
    np.testing.assert_allclose(eps.numpy(), G["eps_threshold"], rtol=1e-6)


def test_ring_evicts_like_the_reference_deque():
    """Batched pushes of ragged sizes leave exactly what 100 deque.append calls leave (golden from ReplayMemory)."""
    ring = ReplayRing(40, 3, "cpu", rngseed=0)
    i = 0
    for n in (1, 7, 32, 40, 20):           # 100 transitions in ragged batches, one of them a full wrap
        ids = torch.arange(i, i + n, dtype=torch.float32)
        ring.push(ids[:, None].expand(n, 3), (torch.arange(i, i + n) % 5), ids[:, None].expand(n, 3) + 0.5, -ids,
                  torch.arange(i, i + n) % 7 == 0)
        i += n
    assert len(ring) == int(G["deque_len"])
    s, a, ns, r, t = ring.ordered()
    np.testing.assert_array_equal(s[:, 0].numpy(), G["deque_state0"])
    np.testing.assert_array_equal(a.numpy(), G["deque_action"])
    np.testing.assert_array_equal(ns[:, 0].numpy(), G["deque_state0"] + 0.5)
    np.testing.assert_array_equal(r.numpy(), -G["deque_state0"])
    # a push larger than the capacity keeps the newest `capacity`
    big = torch.arange(1000, 1100, dtype=torch.float32)
    ring.push(big[:, None].expand(100, 3), torch.zeros(100, dtype=torch.int64), big[:, None].expand(100, 3), big, big < 0)
    ref = deque(range(1000, 1100), maxlen=40)
    np.testing.assert_array_equal(ring.ordered()[0][:, 0].numpy(), np.array(ref, dtype=np.float32))
    assert ring.pushed == 200


def test_ring_sampling_is_uniform_without_replacement():
    ring = ReplayRing(64, 1, "cpu", rngseed=4)
    with pytest.raises(ValueError):
        ring.sample_indices(1)                       # random.sample raises on an empty population too
    v = torch.arange(50, dtype=torch.float32)
    ring.push(v[:, None], torch.zeros(50, dtype=torch.int64), v[:, None], v, v < 0)
    counts = np.zeros(50)
    for _ in range(400):
        idx = ring.sample_indices(10).numpy()
        assert len(set(idx.tolist())) == 10 and idx.max() < 50
        counts[idx] += 1
    assert counts.min() > 40 and counts.max() < 125  # mean 80


def test_select_action_is_epsilon_greedy_per_env():
    ag = VectorDQN(device="cpu", rngseed=2)
    ag.init(golden_net())
    obs = torch.tensor(G["state"])
    greedy = ag.select_action(obs, 0, test=True)
    with torch.no_grad():
        assert torch.equal(greedy, ag.policy_net(obs).max(1)[1])
    late = ag.select_action(obs.repeat(64, 1), torch.full((64 * len(obs),), 10 ** 7))      # eps -> eps_end = 0.05
    frac = (late != greedy.repeat(64)).float().mean().item()
    assert frac < 0.08
    early = ag.select_action(obs.repeat(64, 1), 0)                                        # eps = 0.9, 4/5 of those differ
    assert 0.55 < (early != greedy.repeat(64)).float().mean().item() < 0.85


def test_hard_target_update_fires_on_episode_multiples():
    ag = VectorDQN(device="cpu", rngseed=2)
    ag.params.update({"soft_target_update": False, "target_update": 50})
    ag.init(golden_net())
    with torch.no_grad():
        for p in ag.policy_net.parameters():
            p.add_(1.0)
    ag.update_episode(49, n_finished=3)
    assert not torch.equal(ag.target_net.linear[0].weight, ag.policy_net.linear[0].weight)
    ag.update_episode(52, n_finished=3)              # episodes 50, 51, 52 ended: 50 is a multiple
    assert torch.equal(ag.target_net.linear[0].weight, ag.policy_net.linear[0].weight)


def test_save_state_round_trip():
    ag = VectorDQN(device="cpu", rngseed=2, updates_per_push=1)
    ag.params.update({"batch_size": 8, "min_memory_replay": 8, "memory_replay": 32})
    ag.init(golden_net())
    batch = tuple(torch.tensor(G[k]) for k in ("state", "action", "next_state", "reward", "terminal"))
    ag.update_step(*batch)
    saved = ag.get_save_state()
    ag2 = VectorDQN(device="cpu", rngseed=2)
    ag2.load_save_state(saved)
    assert len(ag2.memory) == len(ag.memory) and ag2.memory.pushed == ag.memory.pushed
    for a, b in zip(ag.memory.ordered(), ag2.memory.ordered()):
        assert torch.equal(a, b)


@pytest.mark.gpu
def test_rollout_on_the_cuda_path():
    """64 envs, 40 vector steps through fused_step with the learner in the loop: transitions are what the envs
    produced (next_obs of step t is obs of step t+1 unless the episode ended), episodes finish and get logged."""
    from gripper_mujoco_b200.mjclass import MjClass
    from gripper_mujoco_b200.model import GripperParams
    from gripper_mujoco_b200.settings import Settings
    st = Settings(); st.debug = False; st.random_seed = 99; st.set_use_noise(True)
    st.mujoco_timestep, st.sim_steps_per_action = 6.857e-3, 30
    mj = MjClass(64, GripperParams(num_segments=5), settings=st, calibrate=False)
    ag = VectorDQN(device=mj.device, rngseed=7, updates_per_push=2)
    ag.params.update({"batch_size": 32, "min_memory_replay": 128, "memory_replay": 4096})
    ag.init(VariableNetwork([mj.get_n_obs(), 64, 64, mj.get_n_actions()], mj.device))
    ro = VectorRollout(mj, ag, max_episode_steps=12)
    w0 = ag.policy_net.linear[0].weight.detach().clone()
    ro.reset_all()
    prev_next, prev_fin = None, None
    for t in range(40):
        obs_before = ro.obs.clone()
        if prev_next is not None:
            keep = ~prev_fin
            assert torch.equal(obs_before[keep], prev_next[keep])
        r, term, trunc = ro.step()
        s, a, ns, rw, tm = ag.memory.ordered()
        assert torch.equal(s[-64:], obs_before) and torch.equal(rw[-64:], r) and torch.equal(tm[-64:], term | trunc)
        assert int(a.min()) >= 0 and int(a.max()) < mj.get_n_actions()
        prev_next, prev_fin = ns[-64:].clone(), (term | trunc)
    out = ro.run(0)
    assert out["episodes"] == 0 and ro.episodes_done >= 64 * 3        # cap 12 => every env finished >= 3 episodes
    assert len(ag.memory) == 40 * 64 and ag.n_updates == 2 * 39        # the first push (64 < 128) cannot optimise
    assert not torch.equal(w0, ag.policy_net.linear[0].weight)
    assert torch.isfinite(ro.returns_sum).item()
    mj.close()
