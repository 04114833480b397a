"""the evaluator's episode loop (/root/reference/drone_evaluator.py:97-186 -- 6 drones, density 0.05,
radius 3, the ten episode seeds, rewards summed in ``rewards.values()`` order) driven end to end through
``dronerl_b200.gym_api`` with the five bundled baseline checkpoints as opponents (``tools/run_evaluator.py``), and the
training loop's call shape of train_torch.py:41-122.  Scores are not expected to equal
tests/torch_tests/test_drone_evaluator.py:5-11 (different rules and RNG): what is asserted is that the loop runs on the
adapter unchanged, is deterministic per seed, and that the scores have the size the rules imply."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'tools'))

needs_models = pytest.mark.skipif(
    not os.path.exists(os.path.join(ROOT, 'baseline', '_ref', 'sample_models', 'dqn-agent-1.safetensors')),
    reason='baseline/_ref/sample_models not installed (__graft_entry__.build() copies it where /root/reference exists)')


@needs_models
def test_evaluator_loop_on_the_adapter():
    import run_evaluator as R
    a = R.evaluate(R.load_qnetwork(os.path.join(R.MODELS_DIR, 'dqn-agent-1.safetensors')), steps=150, seeds=[845, 99])
    b = R.evaluate(R.load_qnetwork(os.path.join(R.MODELS_DIR, 'dqn-agent-1.safetensors')), steps=150, seeds=[845, 99])
    assert a == b                                   # same seeds, same agents -> same episode, same scores
    assert a['n_drones'] == 6 and a['side_size'] == 11
    assert sorted(a['mean_by_agent']) == ['YOU'] + [f'baseline-{i}' for i in range(1, 6)]
    for v in a['mean_by_agent'].values():           # per-step reward is in [-1.1, 1]: bounded by the step count
        assert -1.1 * 150 <= v <= 150
    assert any(v != 0 for v in a['mean_by_agent'].values())
    r = R.evaluate(R.RandomAgent(1), steps=100, seeds=[17])
    assert r['episodes'] == 1 and np.isfinite(r['score'])


@needs_models
def test_checkpoint_networks_have_the_reference_shapes():
    import run_evaluator as R
    x = np.random.default_rng(0).random((3, 7, 7, 6)).astype(np.float32)
    for i in range(1, 6):
        q = R.load_qnetwork(os.path.join(R.MODELS_DIR, f'dqn-agent-{i}.safetensors'))(x)
        assert tuple(q.shape) == (3, 5)
