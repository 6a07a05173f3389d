"""Submap hand-off (eags_raster/handoff.py): the dictionary of
src/entities/submap.py:109-124 survives pack -> move -> unpack bit-exactly, in process (mailbox, two
threads) and between ranks (world_size-2 gloo), and still writes the reference's checkpoint file."""
import os
import threading

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from eags_raster import handoff


def make_submap(i=3, P=1234):
    g = torch.Generator().manual_seed(i)
    return {
        "id": i,
        "gaussian_params": {
            "xyz": torch.randn(P, 3, generator=g), "features_dc": torch.randn(P, 1, 3, generator=g),
            "features_rest": torch.zeros(P, 0, 3), "scaling": torch.randn(P, 3, generator=g),
            "rotation": torch.randn(P, 4, generator=g), "opacity": torch.randn(P, 1, generator=g),
        },
        "Tmc": torch.eye(4, dtype=torch.float64)[None].repeat(5, 1, 1),
        "kf_id": torch.arange(10, 15),
        "T_prev_m": torch.eye(4, dtype=torch.float64),
    }


def assert_same(a, b):
    pa, pb = dict(handoff._walk(a)), dict(handoff._walk(b))
    assert pa.keys() == pb.keys()
    for k in pa:
        assert pa[k].dtype == pb[k].dtype and pa[k].shape == pb[k].shape and torch.equal(pa[k].cpu(), pb[k].cpu()), k
    assert a["id"] == b["id"]


def test_pack_unpack_roundtrip_and_single_block_per_dtype():
    sm = make_submap()
    header, blocks = handoff.pack(sm)
    assert [(dt, dev) for dt, dev, _ in header["groups"]] == [(torch.float32, "cpu"), (torch.float64, "cpu"), (torch.int64, "cpu")]
    assert blocks[0].numel() == 1234 * (3 + 3 + 3 + 4 + 1)
    assert_same(handoff.unpack(header, blocks), sm)


def test_mailbox_between_threads():
    box, sm, got = handoff.SubmapMailbox(), make_submap(7), {}
    t = threading.Thread(target=lambda: got.update(sm=box.get(7, timeout=10)))
    t.start()
    box.put(sm, "cpu")
    t.join()
    assert_same(got["sm"], sm)
    assert box.ids() == [7]
    with pytest.raises(TimeoutError):
        box.get(8, timeout=0.05)
    with pytest.raises(KeyError):
        box.put({"id": 1}, "cpu")


def test_ckpt_layout_matches_reference(tmp_path):
    path = handoff.to_ckpt(make_submap(12), str(tmp_path))
    assert os.path.basename(path) == "000012.ckpt"
    back = torch.load(path, weights_only=False)
    assert tuple(back) == handoff.CKPT_KEYS
    assert_same(back, make_submap(12))


def _worker(rank, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=2)
    if rank == 0:
        handoff.send_submap(make_submap(5), dst=1)
    else:
        assert_same(handoff.recv_submap(0, "cpu"), make_submap(5))
    dist.barrier()
    dist.destroy_process_group()


def test_rank_to_rank_gloo():
    mp.spawn(_worker, args=(29577,), nprocs=2, join=True)


def like_from_model(sm, dev):
    """The device mix Submap.from_model really produces (submap.py:72-84): gaussian_params and Tmc on the
    mapping device, T_prev_m float64 on the CPU (gaussian_slam.py:225-228), kf_id int32 on the CPU."""
    out = dict(sm)
    out["gaussian_params"] = {n: t.to(dev) for n, t in sm["gaussian_params"].items()}
    out["Tmc"] = sm["Tmc"].to(dev)
    out["kf_id"] = sm["kf_id"].int()
    return out


def test_pack_groups_by_dtype_and_device():
    sm = like_from_model(make_submap(4), "meta")      # "meta" stands in for the second device on a CPU box
    groups = {}
    for _, t in handoff._walk(sm):
        groups.setdefault((t.dtype, str(t.device)), 0)
    assert set(groups) == {(torch.float32, "meta"), (torch.float64, "meta"), (torch.float64, "cpu"), (torch.int32, "cpu")}
    header, blocks = handoff.pack(sm)
    assert {(dt, dev) for dt, dev, _ in header["groups"]} == set(groups)
    assert all(b.device.type == dev for b, (_, dev, _) in zip(blocks, header["groups"]))


@pytest.mark.gpu
def test_mailbox_takes_the_mixed_device_dict_of_from_model():
    """Tmc on the GPU + T_prev_m on the CPU used to hit torch.cat's same-device check."""
    sm = make_submap(9, P=20_000)
    src = like_from_model(sm, "cuda:0")
    dst = f"cuda:{torch.cuda.device_count() - 1}"     # another device when the box has one
    box = handoff.SubmapMailbox()
    box.put(src, dst)
    got = box.get(9, pop=True)
    assert all(str(t.device) == dst for _, t in handoff._walk(got))
    ref = dict(sm)
    ref["kf_id"] = sm["kf_id"].int()
    assert_same(got, ref)


@pytest.mark.gpu
def test_mailbox_moves_to_the_gpu():
    box, sm = handoff.SubmapMailbox(), make_submap(2, P=50_000)
    box.put({k: (v if k != "gaussian_params" else {n: t.cuda() for n, t in v.items()}) for k, v in sm.items()}, "cuda")
    got = box.get(2, pop=True)
    assert all(t.is_cuda for _, t in handoff._walk(got))
    assert_same(got, sm)
    assert box.ids() == []
