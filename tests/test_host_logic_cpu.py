"""Host-side logic that needs no GPU: descriptors, group tables, label remap, key packing, and the
two-rank exchange of the sharded push (gloo, world_size 2)."""
import os

import numpy as np
import pytest
import torch

import scaleprotoseg_b200 as sps
from oracle import head_oracle as ho
from scaleprotoseg_b200 import push as spush
from scaleprotoseg_b200.install import _labels_for_scan


def test_config_from_ranges():
    cfg = sps.HeadConfig.from_ranges([(0, 5), (5, 11), (11, 11), (11, 20)], 3, act="linear")
    assert cfg.scale_lo == (0, 5, 11, 11, 20) and cfg.act_kind == sps.ACT_LINEAR
    with pytest.raises(ValueError):
        sps.HeadConfig.from_ranges([(0, 5), (6, 11)], 3)


def test_group_tables_ragged():
    ident = ho.make_identity(24, 3, 4)
    keep = [i for i in range(24) if i not in (0, 5) and ident[i, 1] == 0]  # class 1 loses everything
    ident = ident[keep]
    t = sps.build_group_tables(ident, 3, "cpu")
    assert t.Kc == 2 and t.class_ids == [0, 2]
    members = ho.class_members(ident)
    assert [m.tolist() for m in members] == [m.tolist() for m in t.members_host]
    off = t.class_off.tolist()
    for k, m in enumerate(members):
        assert t.members[off[k]:off[k + 1]].tolist() == m.tolist()
        assert t.proto_slot[m].tolist() == [k] * len(m)
        assert t.proto_col[m].tolist() == list(range(len(m)))


def test_label_remap_matches_reference_onehot():
    K = 5
    t = torch.randint(0, K + 1, (1, 6, 7))
    for void in (0, 2, None):
        lab = _labels_for_scan(t if void is not None else t.clamp(max=K - 1), void)
        tt = t if void is not None else t.clamp(max=K - 1)
        oh = torch.nn.functional.one_hot(tt, K if void is None else K + 1)
        if void is not None:
            oh = torch.cat([oh[..., :void], oh[..., void + 1:]], -1)
        col = torch.where(oh.sum(-1) > 0, oh.argmax(-1) + 1, torch.zeros_like(tt))
        assert torch.equal(lab.to(torch.int64), col)


def test_decode_keys_roundtrip():
    d = torch.tensor([0.0, 1.5, 1e10, 3.25e-5], dtype=torch.float32)
    idx = torch.tensor([0, 7, 0, 123456], dtype=torch.int64)
    keys = (d.view(torch.int32).to(torch.int64) << 32) | idx
    lo, val = spush.decode_keys(keys)
    assert torch.equal(lo, idx) and torch.equal(val, d)
    assert int(keys[2]) >> 32 == spush.ABSENT_BITS
    assert (keys < spush.KEY_INIT).all() and (keys >= 0).all()
    assert torch.equal(torch.argsort(keys), torch.argsort(d, stable=True))  # key order == float order


def _rank_main(rank, world, port, tmp):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    K, P, S, h, w, n = 4, 32, 4, 6, 5, 6
    ident = ho.make_identity(P, K, S)
    ranges = ho.uniform_ranges(P, S)
    protos = torch.rand(P, 64, 1, 1, generator=torch.Generator().manual_seed(1))
    maps = [ho.make_feature_map(i, 256, h, w) for i in range(n)]
    maps[3] = maps[1].clone()  # cross-shard exact tie: the lower global index must win
    labels = [ho.make_labels(i, h, w, K, absent_class=2) for i in range(n)]
    labels[3] = labels[1].clone()
    per = n // world
    lo = rank * per
    # the per-map scan comes from the oracle here (no GPU); the exchange logic is what is under test
    keys = []
    for i in range(lo, lo + per):
        d = ho.scale_l2_convolution(maps[i], protos, ranges)
        idx, val = ho.push_image_min(d, labels[i], ident, K, 0)
        keys.append((val.view(torch.int32).to(torch.int64) << 32) | idx)
    keys = torch.cat(keys)
    best = ((keys & ~0xFFFFFFFF) | (torch.arange(per).unsqueeze(1) + lo)).min(dim=0).values
    dist.all_reduce(best, op=dist.ReduceOp.MIN)
    win = best & 0xFFFFFFFF
    out = torch.zeros(P, 64)
    for p in range(P):
        m = int(win[p]) - lo
        if 0 <= m < per:
            pix = int(keys[m, p] & 0xFFFFFFFF)
            s = p // (P // S)
            out[p] = maps[lo + m].view(S, 64, h * w)[s, :, pix]
    dist.all_reduce(out, op=dist.ReduceOp.SUM)
    if rank == 0:
        ref_win, _, _, ref_new = ho.push_full(maps, labels, protos, ranges, ident, K)
        ok = torch.equal(win, ref_win) and np.array_equal(out.view(P, 64, 1, 1).numpy(), ref_new)
        open(tmp, "w").write("ok" if ok else "mismatch")
    dist.destroy_process_group()


def test_push_two_rank_exchange_gloo(tmp_path):
    import torch.multiprocessing as mp
    out = str(tmp_path / "result.txt")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_rank_main, args=(2, port, out), nprocs=2, join=True)
    assert open(out).read() == "ok"


def test_kld_tables_follow_the_reference_pair_order():
    """Groups = (class, scale) in class-major order, members ascending, pairs (j, k > j) as loss.py:105-127."""
    from scaleprotoseg_b200.kld import KldTables, labels_int16
    ident = ho.make_identity(24, 2, 4)
    keep = [p for p in range(24) if p not in (1, 7, 8)]          # ragged: class 0 keeps one prototype in scale 1
    ranges = [(0, 5), (5, 9), (9, 15), (15, 21)]
    t = KldTables(ident[keep], ranges, "cpu")
    assert t.G == 8 and t.P == 21
    sizes = (t.group_moff[1:] - t.group_moff[:-1]).tolist()
    assert sizes == [2, 1, 3, 3, 3, 3, 3, 3, 0]                  # last = parking group for unassigned prototypes
    assert t.Q == sum(m * (m - 1) // 2 for m in sizes)
    for g in range(t.G):
        ms = t.members[t.group_moff[g]:t.group_moff[g + 1]].tolist()
        assert ms == sorted(ms)
        want = [(ms[i], ms[j]) for i in range(len(ms)) for j in range(i + 1, len(ms))]
        got = list(zip(t.pair_a[t.group_poff[g]:t.group_poff[g + 1]].tolist(),
                       t.pair_b[t.group_poff[g]:t.group_poff[g + 1]].tolist()))
        assert got == want
        for r, p in enumerate(ms):
            assert int(t.proto_group[p]) == g and int(t.proto_local[p]) == r
    lab = labels_int16(torch.tensor([[0, 1, 2, 3, 255, 65537, -1]]), 2)
    assert lab.tolist() == [[0, 1, 2, 0, 0, 0, 0]]


def test_merge_topk_orders_packed_keys_as_unsigned():
    """Per-rank (P,k) lists of packed (distance bits << 32 | map) keys merge to the k smallest; the fill value
    0xFFFF... (negative as int64) sorts last."""
    def key(d, m):
        return (int(np.float32(d).view(np.int32)) << 32) | m
    a = torch.tensor([[key(0.5, 3), key(2.0, 1), -1], [key(1.0, 0), -1, -1]], dtype=torch.int64)
    b = torch.tensor([[key(0.25, 9), key(2.0, 0), key(7.0, 4)], [-1, -1, -1]], dtype=torch.int64)
    pa, pb = torch.arange(6, dtype=torch.int32).view(2, 3), torch.arange(6, 12, dtype=torch.int32).view(2, 3)
    mk, mp = spush.merge_topk([a, b], [pa, pb])
    assert mk[0].tolist() == [key(0.25, 9), key(0.5, 3), key(2.0, 0)] and mp[0].tolist() == [6, 0, 7]
    assert mk[1].tolist() == [key(1.0, 0), -1, -1] and mp[1, 0].item() == 3
