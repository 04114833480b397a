"""Replay ring buffer (dronerl_b200/buffers.py) against the reference's own buffer tests,
/root/reference/tests/jax_tests/test_buffers.py:36-128, restated on torch CPU tensors."""
import torch

from dronerl_b200.buffers import BufferState, ReplayBuffer


def example():
    return {'state': torch.tensor([1.0, 2.0, 3.0]), 'action': torch.tensor([0]), 'reward': torch.tensor([1.0]),
            'next_state': torch.tensor([4.0, 5.0, 6.0]), 'done': torch.tensor([0])}


def test_buffer_initialization():                       # test_buffers.py:36-42
    buf = ReplayBuffer(buffer_size=100, sample_batch_size=10)
    st = buf.init(example())
    assert isinstance(st, BufferState) and st.current_idx == 0 and st.current_size == 0
    for v in st.experiences.values():
        assert v.shape[0] == 100


def test_add_experience():                              # :45-51
    buf = ReplayBuffer(100, 10)
    st = buf.add(buf.init(example()), example())
    assert st.current_idx == 1 and st.current_size == 1
    for k, v in example().items():
        assert torch.all(st.experiences[k][0] == v)


def test_buffer_wrapping():                             # :54-62
    buf = ReplayBuffer(100, 10)
    st = buf.init(example())
    for i in range(105):
        st = buf.add(st, {k: v + i for k, v in example().items()})
    assert st.current_idx == 5 and st.current_size == 100
    assert torch.all(st.experiences['state'][4] == example()['state'] + 104)
    assert torch.all(st.experiences['state'][5] == example()['state'] + 5)


def test_sample_and_can_sample():                       # :65-78
    buf = ReplayBuffer(100, 10)
    st = buf.init(example())
    assert not buf.can_sample(st)
    for i in range(15):
        st = buf.add(st, {k: v + i for k, v in example().items()})
    assert buf.can_sample(st)
    batch = buf.sample(st)
    assert isinstance(batch, dict)
    for v in batch.values():
        assert v.shape[0] == 10
    # only filled slots are ever sampled
    for _ in range(50):
        assert buf.sample(st)['reward'].max() <= 1.0 + 14


def test_add_many_experiences():                        # :82-128
    buf = ReplayBuffer(100, 10)
    ex = example()
    batch = {k: torch.stack([v + i for i in range(5)]) for k, v in ex.items()}
    st = buf.add_many(buf.init(ex), batch)
    assert st.current_idx == 5 and st.current_size == 5
    for i in range(5):
        for k in ex:
            assert torch.all(st.experiences[k][i] == ex[k] + i)
    st = buf.init(ex)
    st.current_idx.fill_(97)                            # wrap: 3 at the end, 2 at the start
    st = buf.add_many(st, batch)
    assert st.current_idx == (97 + 5) % 100
    for i in range(3):
        for k in ex:
            assert torch.all(st.experiences[k][97 + i] == ex[k] + i)
    for i in range(2):
        for k in ex:
            assert torch.all(st.experiences[k][i] == ex[k] + 3 + i)


def test_add_many_more_experiences_than_slots_keeps_the_last_writes():
    """k > buffer_size: the ring indices repeat and the reference's ``.at[indices].set`` keeps the last write per slot
    (jax_impl/buffers.py:57-77) -- deterministically, unlike index_copy_ with duplicate indices."""
    import torch
    from dronerl_b200.buffers import ReplayBuffer
    buf = ReplayBuffer(buffer_size=8, sample_batch_size=4)
    st = buf.init({'x': torch.zeros((), dtype=torch.int64)})
    buf.add_many(st, {'x': torch.arange(3)})                       # current_idx = 3
    buf.add_many(st, {'x': torch.arange(100, 121)})                # 21 experiences into 8 slots
    want = torch.zeros(8, dtype=torch.int64)
    for j, v in enumerate(range(100, 121)):                        # sequential scatter, last write wins
        want[(3 + j) % 8] = v
    assert torch.equal(st.experiences['x'], want)
    assert int(st.current_idx) == (3 + 21) % 8 and int(st.current_size) == 8


def test_sample_uses_integer_draws_and_reaches_every_slot():
    import torch
    from dronerl_b200.buffers import ReplayBuffer
    buf = ReplayBuffer(buffer_size=16, sample_batch_size=4096)
    st = buf.init({'x': torch.zeros((), dtype=torch.int64)})
    buf.add_many(st, {'x': torch.arange(11)})
    g = torch.Generator().manual_seed(0)
    got = buf.sample(st, generator=g)['x']
    assert set(got.tolist()) == set(range(11))                     # only filled slots, all of them
