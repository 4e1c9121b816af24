"""Golden vectors for the vectorised DQN learner (gripper_mujoco_b200/rollout.py) from the reference's own
`Agent_DQN` / `ReplayMemory` (/root/reference/rl/agents/DQN.py), imported here in the build container:

    python tests/golden/make_dqn_golden.py

Writes dqn_learner.npz: initial weights of a small Q-network, a fixed batch of transitions, and for every
optimiser / loss / double-DQN / target-update combination the weights after three reference `update_step`
calls whose replay draw is pinned to that batch; plus the reference's epsilon thresholds and the content of
its deque replay after more pushes than its capacity.
"""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("ref_dqn", "/root/reference/rl/agents/DQN.py")
ref = importlib.util.module_from_spec(spec); spec.loader.exec_module(ref)
spec = importlib.util.spec_from_file_location("ref_networks", "/root/reference/rl/networks.py")
refnet = importlib.util.module_from_spec(spec); spec.loader.exec_module(refnet)

N_OBS, N_ACT, BATCH = 12, 5, 16
LAYERS = [N_OBS, 24, 24, N_ACT]
CASES = [("adamW", "smoothL1Loss", False, True), ("adam", "MSELoss", False, False), ("rmsprop", "Huber", True, True),
         ("adam", "smoothL1Loss", True, False)]


def make_net(seed):
    torch.manual_seed(seed)
    return refnet.VariableNetwork(LAYERS, "cpu")


def main():
    rng = np.random.default_rng(77)
    out = {}
    state = rng.standard_normal((BATCH, N_OBS)).astype(np.float32)
    nxt = rng.standard_normal((BATCH, N_OBS)).astype(np.float32)
    action = rng.integers(0, N_ACT, BATCH).astype(np.int64)
    reward = rng.standard_normal(BATCH).astype(np.float32)
    terminal = rng.random(BATCH) < 0.3
    out.update(state=state, next_state=nxt, action=action, reward=reward, terminal=terminal,
               layers=np.array(LAYERS), n_cases=len(CASES))
    net0 = make_net(5)
    for k, v in net0.state_dict().items():
        out[f"w0/{k}"] = v.numpy().copy()
    for ci, (opt, crit, double, soft) in enumerate(CASES):
        ag = ref.Agent_DQN(device="cpu", rngseed=3)
        ag.params.update({"optimiser": opt, "loss_criterion": crit, "use_double_dqn": double, "soft_target_update": soft,
                          "batch_size": BATCH, "min_memory_replay": BATCH, "memory_replay": 64, "learning_rate": 1e-2,
                          "grad_clamp_value": 0.5, "gamma": 0.97})
        ag.init(make_net(5))
        T = ref.Agent_DQN.Transition
        fixed = T(tuple(torch.tensor(state[i:i + 1]) for i in range(BATCH)),
                  tuple(torch.tensor([[action[i]]]) for i in range(BATCH)),
                  tuple(torch.tensor(nxt[i:i + 1]) for i in range(BATCH)),
                  tuple(torch.tensor([reward[i]]) for i in range(BATCH)),
                  tuple(torch.tensor([terminal[i]]) for i in range(BATCH)))
        for i in range(BATCH):       # fill the memory so the length gates pass
            ag.memory.push(fixed.state[i], fixed.action[i], fixed.next_state[i], fixed.reward[i], fixed.terminal[i])
        ag.memory.batch = lambda bs, fixed=fixed: fixed
        for _ in range(3):
            ag.update_step(fixed.state[0], fixed.action[0], fixed.next_state[0], fixed.reward[0], fixed.terminal[0], None)
        out[f"case{ci}/cfg"] = np.array([opt, crit, str(int(double)), str(int(soft))])
        for k, v in ag.policy_net.state_dict().items():
            out[f"case{ci}/policy/{k}"] = v.numpy().copy()
        for k, v in ag.target_net.state_dict().items():
            out[f"case{ci}/target/{k}"] = v.numpy().copy()
    # epsilon schedule (rl/agents/DQN.py:196-201)
    ag = ref.Agent_DQN(device="cpu", rngseed=3)
    decay = np.array([0, 1, 10, 250, 1000, 5000, 120000], dtype=np.int64)
    p = ag.params
    out["eps_decay_num"] = decay
    out["eps_threshold"] = np.array([p.eps_end + (p.eps_start - p.eps_end) * np.exp(-1. * d / float(p.eps_decay)) for d in decay])
    # deque replay content after 2.5 x capacity pushes
    mem = ref.ReplayMemory(40, "cpu", ref.Agent_DQN.Transition)
    for i in range(100):
        mem.push(torch.full((1, 3), float(i)), torch.tensor([[i % 5]]), torch.full((1, 3), float(i) + 0.5),
                 torch.tensor([float(-i)]), torch.tensor([i % 7 == 0]))
    out["deque_state0"] = np.array([t.state[0, 0].item() for t in mem.memory])
    out["deque_action"] = np.array([t.action.item() for t in mem.memory])
    out["deque_len"] = len(mem)
    np.savez_compressed(os.path.join(HERE, "dqn_learner.npz"), **out)
    print("wrote dqn_learner.npz", len(out), "arrays")


if __name__ == "__main__":
    sys.exit(main())
