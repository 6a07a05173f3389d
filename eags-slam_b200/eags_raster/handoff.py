"""Submap hand-off without the disk round trip.

The reference passes a finished submap from the mapping loop to the loop-closure worker through a
file: ``Submap.save`` writes ``{id, gaussian_params, Tmc, kf_id, T_prev_m}`` with ``torch.save``
(src/entities/submap.py:109-124) and ``Loop_closure.submap_loader`` reads it back on the LC device
(src/entities/lc.py:126-171).  With one submap per GPU for parallel loop closure that file is pure
latency.  Here the same dictionary travels

* inside one process (the reference's layout: main thread on device 0, LC thread on ``lc.device``)
  through :class:`SubmapMailbox`: all tensors are packed into ONE contiguous block per (dtype, source
  device) -- the reference's dictionary mixes GPU and CPU tensors -- each moved with one copy
  (device-to-device over NVLink for the Gaussian parameters) and re-viewed on the destination -- no
  serialisation, no host bounce;
* between ranks (one process per GPU) through :func:`send_submap` / :func:`recv_submap`:
  a small pickled header plus one ``torch.distributed`` send per block (NCCL P2P on GPUs, gloo
  in the CPU tests).

The dictionary layout is exactly the checkpoint's, so ``to_ckpt`` still writes a file the
reference's ``Submap.load`` understands (crash recovery / final export).
"""
from __future__ import annotations

import os
import threading
from typing import Any, Dict, List, Optional, Tuple

import torch

CKPT_KEYS = ("id", "gaussian_params", "Tmc", "kf_id", "T_prev_m")


def _walk(obj: Any, path: Tuple = ()):
    """Yield (path, tensor) for every tensor in a nested dict / list / tuple."""
    if torch.is_tensor(obj):
        yield path, obj
    elif isinstance(obj, dict):
        for k, v in obj.items():
            yield from _walk(v, path + (k,))
    elif isinstance(obj, (list, tuple)):
        for i, v in enumerate(obj):
            yield from _walk(v, path + (i,))


def _skeleton(obj: Any):
    """The structure with every tensor replaced by None (picklable header)."""
    if torch.is_tensor(obj):
        return None
    if isinstance(obj, dict):
        return {k: _skeleton(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_skeleton(v) for v in obj)
    return obj


def _set(root: Any, path: Tuple, value: Any):
    if isinstance(root, tuple):
        raise TypeError("tuples holding tensors are not supported in a submap dictionary")
    for k in path[:-1]:
        root = root[k]
    root[path[-1]] = value


def pack(submap: Dict) -> Tuple[Dict, List[torch.Tensor]]:
    """-> (header, blocks).  Tensors are grouped by (dtype, device) -- the reference's dictionary mixes
    devices: ``gaussian_params`` and ``Tmc`` live on the mapping GPU while ``T_prev_m`` (float64) and
    ``kf_id`` (int32) stay on the CPU (submap.py:72-84, gaussian_slam.py:225-228) -- and every group is
    ONE contiguous 1-D block on its source device.  The header records, for every tensor, its path,
    group, shape and offset inside the group's block."""
    entries, chunks, groups = [], [], {}
    for path, t in _walk(submap):
        key = (t.dtype, str(t.device))
        gid = groups.setdefault(key, len(groups))
        if gid == len(chunks):
            chunks.append([])
        lst = chunks[gid]
        off = sum(c.numel() for c in lst)
        entries.append((path, gid, tuple(t.shape), off))
        lst.append(t.detach().reshape(-1))
    blocks = [torch.cat(lst) if len(lst) > 1 else lst[0].clone() for lst in chunks]
    header = {"skeleton": _skeleton(submap), "entries": entries,
              "groups": [(dt, dev, blocks[gid].numel()) for (dt, dev), gid in groups.items()]}
    return header, blocks


def unpack(header: Dict, blocks: List[torch.Tensor]) -> Dict:
    import copy
    out = copy.deepcopy(header["skeleton"])
    for path, gid, shape, off in header["entries"]:
        n = 1
        for s in shape:
            n *= s
        _set(out, path, blocks[gid][off:off + n].view(shape))
    return out


def check_layout(submap: Dict) -> None:
    missing = [k for k in CKPT_KEYS if k not in submap]
    if missing:
        raise KeyError(f"submap dictionary lacks {missing} (submap.py:109-124 layout)")


class SubmapMailbox:
    """In-process hand-off between the mapping thread and the loop-closure thread."""

    def __init__(self):
        self._lock = threading.Condition()
        self._box: Dict[int, Dict] = {}

    def put(self, submap: Dict, device) -> None:
        """Move ``submap`` to ``device`` (one copy per (dtype, source device) block) and publish it under its id."""
        check_layout(submap)
        device = torch.device(device)
        header, blocks = pack(submap)
        # everything ends up on ``device``, as ``torch.load(path, map_location=device)`` leaves it
        # (submap.py:28); one copy per (dtype, source device) group
        moved = [b.to(device, non_blocking=True) for b in blocks]
        if device.type == "cuda":
            # the consumer runs on another thread / stream: hand over only completed copies
            torch.cuda.current_stream(device).synchronize()
        ready = unpack(header, moved)
        with self._lock:
            self._box[int(submap["id"])] = ready
            self._lock.notify_all()

    def get(self, submap_id: int, timeout: Optional[float] = None, pop: bool = False) -> Dict:
        with self._lock:
            if not self._lock.wait_for(lambda: submap_id in self._box, timeout):
                raise TimeoutError(f"submap {submap_id} was not handed off within {timeout} s")
            return self._box.pop(submap_id) if pop else self._box[submap_id]

    def ids(self) -> List[int]:
        with self._lock:
            return sorted(self._box)


def _wire_device(group=None) -> torch.device:
    """Where tensors must live to cross the process group: the current CUDA device under NCCL, else CPU."""
    import torch.distributed as dist
    return torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" \
        else torch.device("cpu")


def send_submap(submap: Dict, dst: int, group=None) -> None:
    """Rank-to-rank hand-off (parallel loop closure, one submap per GPU): a pickled header, then one
    send per (dtype, source device) block (NCCL P2P over NVLink for the GPU-resident blocks)."""
    import torch.distributed as dist
    check_layout(submap)
    header, blocks = pack(submap)
    wire = _wire_device(group)
    dist.send_object_list([header], dst=dst, group=group)
    for b in blocks:
        dist.send(b.to(wire).contiguous(), dst=dst, group=group)


def recv_submap(src: int, device, group=None) -> Dict:
    import torch.distributed as dist
    box = [None]
    dist.recv_object_list(box, src=src, group=group)
    header = box[0]
    wire, device = _wire_device(group), torch.device(device)
    blocks = []
    for dt, _dev, n in header["groups"]:
        buf = torch.empty(n, dtype=dt, device=wire)
        dist.recv(buf, src=src, group=group)
        blocks.append(buf.to(device))
    return unpack(header, blocks)


def to_ckpt(submap: Dict, directory: str) -> str:
    """Write the reference's checkpoint file (``{id:06d}.ckpt``, submap.py:122-124)."""
    check_layout(submap)
    os.makedirs(directory, exist_ok=True)
    path = os.path.join(directory, f"{str(int(submap['id'])).zfill(6)}.ckpt")
    torch.save({k: submap[k] for k in CKPT_KEYS}, path)
    return path
