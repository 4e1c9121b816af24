"""BatchedMjEnv (gripper_mujoco_b200/mjenv.py): the MjEnv protocol of rl/env/MjEnv.py over a batch."""
import numpy as np
import pytest
import torch

from gripper_mujoco_b200.objects import build_object_set, task_split


def test_task_split_is_a_partition_into_files_of_20():
    """random_object_split (description/mujoco/xml_script.py:603-660): every object in exactly one task file,
    20 per file, deterministic for a seed; the first 5 files (100 objects) are MjEnv's test set."""
    n = len(build_object_set())
    files = task_split(n)
    assert len(files) == -(-n // 20) and all(len(f) == 20 for f in files[:-1])
    allidx = np.concatenate(files)
    assert sorted(allidx.tolist()) == list(range(n))
    assert all(np.array_equal(a, b) for a, b in zip(files, task_split(n)))
    assert not np.array_equal(np.concatenate(task_split(n, seed=7)), allidx)
    assert np.array_equal(np.concatenate(task_split(n, shuffle=False)), np.arange(n))


def _env(n, **kw):
    from gripper_mujoco_b200.mjenv import BatchedMjEnv
    from gripper_mujoco_b200.settings import Settings
    st = Settings(); st.debug = False; st.set_use_noise(True)
    st.mujoco_timestep, st.sim_steps_per_action = 6.857e-3, 30
    return BatchedMjEnv(n, num_segments=5, seed=11, settings=st, calibrate=False, **kw)


@pytest.mark.gpu
def test_env_protocol_truncation_and_done_guard():
    env = _env(6)
    env.params.max_episode_steps = 4
    obs = env.reset()
    assert obs.shape == (6, env.n_obs) and env.n_actions == env.mj.get_n_actions()
    names = env.mj.get_current_object_name()
    train = set(np.concatenate(env.task_files[env.testing_xmls:]).tolist())
    by_name = {s.name: i for i, s in enumerate(env.objects)}
    assert all(by_name[nm] in train for nm in names)            # training draws never touch the test files
    for t in range(4):
        obs, r, term, trunc, info = env.step(torch.full((6,), 2, dtype=torch.int32))
        assert r.shape == (6,) and term.dtype == torch.bool and trunc.dtype == torch.bool
        assert bool(trunc.all()) == (t == 3) and not bool((term & trunc).any())
    assert torch.allclose(env.cumulative_reward, env.cumulative_reward) and bool(env.is_done.all())
    with pytest.raises(RuntimeError, match="done=true"):
        env.step(torch.zeros(6, dtype=torch.int32))
    env.reset(env.is_done)
    assert int(env.current_step.sum()) == 0 and float(env.cumulative_reward.abs().sum()) == 0.0
    env.step(torch.zeros(6, dtype=torch.int32))
    env.close()


@pytest.mark.gpu
def test_env_seed_reproduces_objects_and_observations():
    a, b = _env(8), _env(8)
    oa, ob = a.reset(), b.reset()
    assert a.mj.get_current_object_name() == b.mj.get_current_object_name()
    assert torch.equal(oa, ob)
    act = torch.arange(8, dtype=torch.int32) % a.n_actions
    for _ in range(3):
        ra, rb = a.step(act), b.step(act)
        assert torch.equal(ra[0], rb[0]) and torch.equal(ra[1], rb[1])
    a.close(); b.close()


@pytest.mark.gpu
def test_env_test_protocol_walks_objects_times_trials():
    """4 test objects x 2 trials dealt to 3 envs in the reference's order (file, object, trial); every trial is
    reported once with its object name and return, then test_completed is set."""
    env = _env(3)
    p = env.params
    p.max_episode_steps, p.test_objects, p.test_obj_per_file, p.test_trials_per_object = 3, 4, 2, 2
    env.task_files = task_split(len(env.objects), obj_per_task=2)
    env.testing_xmls = 2
    data = env.run_test(lambda obs: torch.full((3,), 2, dtype=torch.int32, device=obs.device))
    assert env.test_completed and not env.test_in_progress
    assert len(data) == 8
    got = sorted((d.xml_file, d.obj_idx, d.obj_trial) for d in data)
    assert got == [(f, o, t) for f in range(2) for o in range(2) for t in range(2)]
    for d in data:
        assert d.object_name == env.objects[int(env.task_files[d.xml_file][d.obj_idx])].name
        assert d.steps == 3 and np.isfinite(d.reward) and d.object_category == d.object_name.split("_")[0]
    env.close()
