#!/usr/bin/env python
"""The AIcrowd evaluator's episode loop (/root/reference/drone_evaluator.py:97-186) driven through the drop-in dict
adapter ``dronerl_b200.gym_api`` instead of the reference's torch env.

Same protocol: the five bundled baseline agents plus the submission ("YOU") share one env -- 6 drones, drone_density
0.05 (side 11), WindowedGridView radius 3, seeds [845, 99, 65, 96, 85, 39, 51, 17, 52, 35], 1000 steps per episode,
greedy ``argmax Q`` actions in sorted-agent-name order, per-step rewards summed in ``rewards.values()`` order, score =
mean over episodes (secondary = std).  Video frames / AIcrowd plumbing are left out.  The agents are rebuilt from the
safetensors checkpoints with plain torch modules following the checkpoint format of torch_impl/agents/dqn.py:47-224
(``network.dense_k`` / ``network.conv2d_k`` tensors; ``torch_impl.agents.dqn`` itself imports matplotlib, absent here).

The env behind the adapter follows the JAX rules and the threefry stream (documented in gym_api.py), so the scores are
NOT expected to equal tests/torch_tests/test_drone_evaluator.py:5-11 (MT19937, sequential movers): report the
distribution.  usage: python tools/run_evaluator.py [--steps 1000] [--submission N|random]
"""
import argparse
import ast
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

EPISODE_SEEDS = [845, 99, 65, 96, 85, 39, 51, 17, 52, 35]          # drone_evaluator.py:28
MODELS_DIR = os.path.join(ROOT, 'baseline', '_ref', 'sample_models')


def load_qnetwork(path):
    """A checkpoint of torch_impl/agents/dqn.py (dense or conv) as ``states [B,7,7,6] -> Q [B,5]`` on the CPU."""
    from safetensors import safe_open
    with safe_open(path, framework='pt', device='cpu') as f:
        meta = f.metadata()
        sd = {k: f.get_tensor(k) for k in f.keys()}
    kind = meta['network_type']
    n_dense = len(ast.literal_eval(meta['dense_layers'])) + 1
    n_conv = len(ast.literal_eval(meta['conv_layers'])) if kind == 'conv' else 0
    convs = ast.literal_eval(meta['conv_layers']) if kind == 'conv' else ()

    def forward(states):
        x = torch.as_tensor(np.asarray(states), dtype=torch.float32)
        if kind == 'conv':
            x = x.permute(0, 3, 1, 2)                              # dqn.py:153-157
            for i in range(n_conv):
                x = torch.relu(torch.nn.functional.conv2d(x, sd[f'network.conv2d_{i + 1}.weight'], sd[f'network.conv2d_{i + 1}.bias'],
                                                          stride=convs[i].get('stride', 1), padding=convs[i].get('padding', 0)))
            x = x.flatten(1)
        else:
            x = x.reshape(x.shape[0], -1)                          # dqn.py:80-82
        for i in range(n_dense):
            if i > 0:
                x = torch.relu(x)
            x = torch.nn.functional.linear(x, sd[f'network.dense_{i + 1}.weight'], sd[f'network.dense_{i + 1}.bias'])
        return x
    return forward


class RandomAgent:
    """Uniform random actions with a private generator: Q-values whose argmax is the drawn action."""

    def __init__(self, seed=0):
        self.rng = np.random.default_rng(seed)

    def __call__(self, states):
        q = torch.zeros((len(states), 5))
        q[torch.arange(len(states)), torch.as_tensor(self.rng.integers(0, 5, size=len(states)))] = 1.0
        return q


def evaluate(submission, steps=1000, seeds=EPISODE_SEEDS, models_dir=MODELS_DIR, device=None):
    """``submission``: a callable ``states -> Q`` (the "YOU" agent).  Returns the evaluator's result dict."""
    from dronerl_b200.gym_api import DeliveryDrones, WindowedGridView
    agents = {f'baseline-{i}': load_qnetwork(os.path.join(models_dir, f'dqn-agent-{i}.safetensors')) for i in range(1, 6)}
    agents['YOU'] = submission
    names = sorted(agents)                                          # drone_evaluator.py:54-56: ids by sorted name
    overall = []
    for seed in seeds:
        env_params = {'charge_reward': -0.1, 'crash_reward': -1, 'delivery_reward': 1, 'charge': 20, 'discharge': 10,
                      'drone_density': 0.05, 'dropzones_factor': 2, 'n_drones': len(names), 'packets_factor': 3,
                      'pickup_reward': 0, 'rgb_render_rescale': 1.0, 'skyscrapers_factor': 3, 'stations_factor': 2}
        env = WindowedGridView(DeliveryDrones(env_params, device=device), radius=3)     # :130
        env.reset(seed=seed)                                        # set_seed(env, episode_seed)  (rl_helpers.py:12-18)
        env.action_space.seed(seed)
        state = env.reset()                                         # :144
        scores = np.zeros(len(names))
        with torch.no_grad():
            for _ in range(steps):
                actions = {idx: int(agents[name]([state[idx]])[0].argmax().item()) for idx, name in enumerate(names)}   # :153-167
                state, rewards, _, _, _ = env.step(actions)         # :170
                scores += np.array(list(rewards.values()))          # :173-175
        overall.append(scores)
    overall = np.array(overall)
    you = names.index('YOU')
    return {'score': float(overall.mean(axis=0)[you]), 'score_secondary': float(overall.std(axis=0)[you]),
            'mean_by_agent': dict(zip(names, overall.mean(axis=0).round(3).tolist())),
            'std_by_agent': dict(zip(names, overall.std(axis=0).round(3).tolist())),
            'episodes': len(seeds), 'steps': steps, 'side_size': 11, 'n_drones': len(names)}


if __name__ == '__main__':
    ap = argparse.ArgumentParser()
    ap.add_argument('--steps', type=int, default=1000)
    ap.add_argument('--submission', default='1', help='1..5: a bundled baseline as the submission; "random": a random agent')
    a = ap.parse_args()
    sub = RandomAgent(0) if a.submission == 'random' else load_qnetwork(os.path.join(MODELS_DIR, f'dqn-agent-{a.submission}.safetensors'))
    out = evaluate(sub, steps=a.steps)
    out['submission'] = a.submission
    print(json.dumps(out))
