"""Bucketed forward (stage 1 scatters the keys into per-tile buckets, no duplicate kernel): results
must equal the exact path bit for bit -- the per-tile sort is deterministic, so ranges, point_list,
sorted keys, n_contrib AND the rendered images are identical -- and an undersized bucket must fall
back to the exact path transparently."""
import pytest
import torch

from eags_raster import rasterize as rz
from eags_raster import scene, state_array
from helpers import cuda_forward

pytestmark = pytest.mark.gpu
NAMES = ("ranges", "point_list", "point_list_keys", "n_contrib", "final_T", "tiles_touched")


def _run(cam, sc):
    st, d, out = cuda_forward(cam, sc)
    state = out[5]
    K = state.num_rendered
    arrays = {n: state_array(state, n) for n in NAMES}
    arrays["point_list"] = arrays["point_list"][:K]
    arrays["point_list_keys"] = arrays["point_list_keys"][:K]
    return out, arrays


@pytest.mark.parametrize("name,P", [("ragged", 3000), ("c1", 50_000)])
def test_bucketed_equals_exact(name, P):
    sc = scene.make_scene(name, P=P, seed=2)
    cam = scene.workload_camera(name)
    key = (torch.cuda.current_device(), cam.W, cam.H)
    try:
        rz.set_bucketed(False)
        ref_out, ref = _run(cam, sc)
        rz.set_bucketed(True)
        rz._tile_cap.pop(key, None)
        _run(cam, sc)                                   # learns the bucket size (exact path)
        assert rz._tile_cap[key] >= int((ref["ranges"][:, 1] - ref["ranges"][:, 0]).max())
        out, got = _run(cam, sc)                        # bucketed
        for n in NAMES:
            assert torch.equal(got[n], ref[n]), n
        for a, b in zip(out[:4], ref_out[:4]):
            assert torch.equal(a, b)
        # undersized buckets: device flag -> exact re-issue, same results, bound raised
        rz._tile_cap[key] = 64
        out, got = _run(cam, sc)
        for n in NAMES:
            assert torch.equal(got[n], ref[n]), n
        for a, b in zip(out[:4], ref_out[:4]):
            assert torch.equal(a, b)
        assert rz._tile_cap[key] > 64
    finally:
        rz.set_bucketed(True)


def test_status_block_reaches_pinned_and_pageable_host_memory(monkeypatch):
    """The forward's status block (K, largest tile, overflow flag) is stored to the caller's buffer by the tile
    scan itself when that buffer is pinned host memory, and copied by a cudaMemcpyAsync node when it is pageable:
    both must report the same K and render the same image."""
    sc = scene.make_scene("ragged", P=3000, seed=4)
    cam = scene.workload_camera("ragged")
    rz.set_bucketed(True)
    ref_out, _ = _run(cam, sc)
    ref_out, _ = _run(cam, sc)                           # second call: bucketed, pinned status (the default)
    pageable = torch.zeros(8, dtype=torch.int64)
    assert not pageable.is_pinned()
    monkeypatch.setattr(rz, "_pinned_status", lambda: pageable)
    pageable.fill_(-1)
    out, _ = _run(cam, sc)
    assert out[5].num_rendered == ref_out[5].num_rendered > 0 and int(pageable[0]) == out[5].num_rendered
    assert torch.equal(out[0], ref_out[0]) and torch.equal(out[1], ref_out[1])
