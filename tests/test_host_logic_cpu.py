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


def _dropin_rank_main(rank, world, port, tmp, ann_dir):
    """The INSTALLED push drop-ins (install.make_push_functions) on `world` gloo ranks.  The three CUDA entry points they
    call are replaced by oracle-based CPU stand-ins (no GPU here): what is under test is the sharding of dataset.img_ids,
    the batching of same-size maps, the min all-reduce of packed keys, the all-gather of the per-image index lists and
    the sum all-reduce of the gathered winners - against the oracle's single-process push."""
    import types
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from scaleprotoseg_b200 import push as sp
    from scaleprotoseg_b200.install import make_push_functions
    from standin_model import Stage1Standin
    K, P, S, n = 4, 32, 4, 7
    sizes = [(6, 5)] * 3 + [(4, 7)] * 2 + [(6, 5)] * 2            # a change of map size forces a batch flush
    torch.manual_seed(3)
    model = Stage1Standin(P, K)
    xs = [torch.randn(256, h, w, generator=torch.Generator().manual_seed(50 + i)) for i, (h, w) in enumerate(sizes)]
    xs[5] = xs[0].clone()          # cross-shard exact tie: the lower global index must win
    labels = [ho.make_labels(i, h, w, K, absent_class=2) for i, (h, w) in enumerate(sizes)]
    labels[5] = labels[0].clone()
    ident, ranges = model.prototype_class_identity, model._ranges()

    def scan_maps(z, lab, protos, pc, cfg, keys_out=None):
        rows = []
        for i in range(z.shape[0]):
            d = ho.scale_l2_convolution(z[i:i + 1], protos, ranges)
            idx, val = ho.push_image_min(d, lab[i].long(), ident, K, 0)
            rows.append((val.view(torch.int32).to(torch.int64) << 32) | idx)
        keys = torch.cat(rows)
        if keys_out is not None:
            keys_out.copy_(keys)
        return keys

    def reduce_keys(keys, map_base, best=None):
        k = (keys & ~0xFFFFFFFF) | (torch.arange(keys.shape[0]).unsqueeze(1) + map_base)
        return k.min(dim=0).values

    def gather_winners(z, keys, best, map_base, proto_scale, cfg, out=None):
        _, C, H, W = z.shape
        for p in range(keys.shape[1]):
            if int(best[p] & 0xFFFFFFFF) == map_base:
                pix = int(keys[0, p] & 0xFFFFFFFF)
                out[p] = z[0].view(S, C // S, H * W)[int(proto_scale[p]), :, pix]
        return out

    sp.scan_maps, sp.reduce_keys, sp.gather_winners = scan_maps, reduce_keys, gather_winners
    ids = [f"i{i}" for i in range(n)]
    ds = types.SimpleNamespace(img_ids=ids, annotations_dir=ann_dir, convert_targets=None, image_margin_size=0)
    pm = types.SimpleNamespace(to_normalized_tensor=lambda t: t, resize_label=lambda t, size: torch.from_numpy(np.asarray(t)).long())
    maps = dict(zip(ids, xs))
    _, min_across_dataset, global_min = make_push_functions(pm, load_image=lambda d, iid: maps[iid])
    win, list_idx = min_across_dataset(ds, model, K, void_class=0, device="cpu")
    patches = global_min(win, list_idx, ds, model, device="cpu")
    ref_win, ref_idx, _, ref_new = ho.push_full([torch.sigmoid(x).unsqueeze(0) for x in xs], labels, model.prototype_vectors.detach(),
                                                ranges, ident, K)
    ok = torch.equal(win, ref_win) and all(torch.equal(a, b) for a, b in zip(list_idx, ref_idx))
    ok = ok and np.array_equal(np.reshape(patches, ref_new.shape), ref_new)
    open(tmp + f".{rank}", "w").write("ok" if ok else "mismatch")
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_push_dropins_sharded_gloo(tmp_path, world):
    """Row e' of the scope table: the reference's unchanged push_prototypes_multiscale calls min_across_dataset / global_min
    as module globals (:286,:291); with torch.distributed initialised the installed versions shard the dataset."""
    import torch.multiprocessing as mp
    n = 7
    sizes = [(6, 5)] * 3 + [(4, 7)] * 2 + [(6, 5)] * 2
    for i, (h, w) in enumerate(sizes):
        lab = ho.make_labels(i if i != 5 else 0, h, w, 4, absent_class=2)
        np.save(str(tmp_path / f"i{i}.npy"), lab.numpy())
    out = str(tmp_path / "result")
    port = 29500 + ((os.getpid() + 17 * world) % 2000)
    mp.spawn(_dropin_rank_main, args=(world, port, out, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert open(out + f".{r}").read() == "ok", f"rank {r}"


def test_shard_bounds_cover_the_dataset_contiguously():
    from scaleprotoseg_b200.install import shard_bounds
    for n in (0, 1, 7, 8, 3000):
        for world in (1, 2, 3, 8):
            cuts = [shard_bounds(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:])) and all(lo <= hi for lo, hi in cuts)


def test_install_on_the_real_reference_classes():
    """install() on the UNMODIFIED reference (skipped where /root/reference is absent, i.e. on the GPU box): every patched
    method and push global exists there, keeps its parameter names, and uninstall() puts the originals back."""
    import inspect
    import sys
    ref = "/root/reference"
    if not os.path.isdir(os.path.join(ref, "segmentation")):
        pytest.skip("reference checkout not present")
    for k in ["DATA_PATH_CITY", "DATA_PATH_PASCAL", "DATA_PATH_ADE", "DATA_PATH_EM", "DATA_PATH_COCO", "LOG_DIR"]:
        os.environ.setdefault(k, "/tmp/sps_ref_env")
    os.makedirs("/tmp/sps_ref_env", exist_ok=True)
    for p in (ref, os.path.join(os.path.dirname(os.path.abspath(__file__)), "_stubs")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import scaleprotoseg_b200 as sps
    from segmentation.model import loss as ml
    from segmentation.model import model as m1
    from segmentation.model import model_multiscale as mm
    from segmentation.model import model_multiscale_group as mg
    import segmentation.push_multiscale_optimization as pm
    import segmentation.push_optimization as pm1
    targets = [(mm.PPNetMultiScale, ["_scale_l2_convolution", "prototype_distances", "push_forward", "forward_from_conv_features"]),
               (mg.PPNetMultiScale, ["_scale_l2_convolution", "push_forward", "compute_group", "run_last_layer", "forward_from_conv_features"]),
               (m1.PPNet, ["_l2_convolution", "prototype_distances", "push_forward", "forward_from_conv_features"]),
               (ml.KLDLoss, ["forward"]), (ml.KLDLossGroup, ["forward"]),
               (pm, ["compute_distances", "min_across_dataset", "global_min"]),
               (pm1, ["compute_distances", "min_across_dataset", "global_min"])]
    before = {(id(o), n): getattr(o, n) for o, names in targets for n in names}
    sig = {k: list(inspect.signature(v).parameters) for k, v in before.items()}
    sps.install()
    try:
        for o, names in targets:
            for n in names:
                new = getattr(o, n)
                assert new is not before[(id(o), n)], f"{o.__name__}.{n} was not replaced"
                assert list(inspect.signature(new).parameters) == sig[(id(o), n)], f"{o.__name__}.{n}: parameter names differ"
        # the push driver resolves its three helpers as module globals at call time
        assert pm.push_prototypes_multiscale.__globals__["min_across_dataset"] is pm.min_across_dataset
        assert pm1.push_prototypes.__globals__["global_min"] is pm1.global_min
    finally:
        sps.uninstall()
    for o, names in targets:
        for n in names:
            assert getattr(o, n) is before[(id(o), n)], f"{o.__name__}.{n} was not restored"


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
