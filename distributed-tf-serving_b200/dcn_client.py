"""Host-side mirror of the reference client's interface on top of the engine.

Same names, argument meaning and error behaviour as
``/root/reference/src/main/java/tensorflow/serving/DCNClient.java`` (``DC``), so that a parity
test reads like a test of the reference: ``partitionList`` / ``getFeatureIds`` / ``getFeatureWts``
/ ``sendRequest`` / ``processARequest``.  What changed underneath: a "channel" is a GPU of the
engine, ``sendRequest`` is a C-ABI call instead of a protobuf RPC, and the fan-out over hosts
is the engine's fan-out over devices.
"""
from __future__ import annotations

from concurrent.futures import Future, ThreadPoolExecutor
from typing import List, Sequence

import numpy as np

from . import binding

FIELD_NUM = 43            # DC:25
candidateNum = 1500       # DC:29
requestNum = 1000         # DC:30
concurrentNum = 6         # DC:31
modelName = "DCN"         # DC:33 (kept for the record; the engine addresses its model by weight file)
modelSignature = "serving_default"   # DC:34
outputKey = "prediction_node"        # DC:35

threadPool = ThreadPoolExecutor(16)  # DC:42


def partitionList(lst: Sequence, partNum: int) -> List:
    """DC:46-55, element-wise like the reference; bounds come from the C ABI."""
    return [lst[a:b] for a, b in binding.partition_bounds(len(lst), partNum)]


def getFeatureIds(n_hosts: int, candidate_num: int = candidateNum, field_num: int = FIELD_NUM) -> List[np.ndarray]:
    """DC:57-65."""
    flat = np.tile(np.arange(1, field_num + 1, dtype=np.int64), candidate_num)
    return partitionList(flat, n_hosts)


def getFeatureWts(n_hosts: int, candidate_num: int = candidateNum, field_num: int = FIELD_NUM) -> List[np.ndarray]:
    """DC:67-74."""
    return partitionList(np.ones(candidate_num * field_num, dtype=np.float32), n_hosts)


class Channel:
    """Stands in for a gRPC ManagedChannel (DC:118-125): one GPU of one engine."""

    def __init__(self, engine: "binding.Engine"):
        self.engine = engine


def getChannels(device_ids: Sequence[int], weight_file: str, max_candidates: int = 16384) -> List[Channel]:
    """DC:118-125: one channel per host -> one single-GPU engine per device, every one holding
    the full model (the reference's servers are replicas)."""
    chans = []
    for d in device_ids:
        e = binding.Engine(device_ids=(d,), max_candidates=max_candidates)
        e.load_weights(weight_file)
        chans.append(Channel(e))
    return chans


def shutDownChannels(channelList: Sequence[Channel]) -> None:
    """DC:127-135 (the reference only sleeps there; this really releases the devices)."""
    for c in channelList:
        c.engine.close()


def sendRequest(channel: Channel, featureIds: np.ndarray, featureWts: np.ndarray, field_num: int = None) -> np.ndarray:
    """DC:91-115: declare shape [len/F, F], ship both tensors, return the model's scores
    (``outputs["prediction_node"].float_val``, DC:162).  A value count that disagrees with the
    declared shape raises, as the RPC would (INVALID_ARGUMENT)."""
    F = field_num or channel.engine.num_fields
    ids = np.ascontiguousarray(featureIds, dtype=np.int64)
    wts = np.ascontiguousarray(featureWts, dtype=np.float32)
    rows = ids.size // F                                     # DC:97 (integer division)
    if rows * F != ids.size or wts.size != ids.size:
        raise ValueError("declared shape [%d, %d] disagrees with %d ids / %d weights" % (rows, F, ids.size, wts.size))
    return channel.engine.predict(ids, wts)


def processARequest(channelList: Sequence[Channel], featIdsList: Sequence[np.ndarray],
                    featWtsList: Sequence[np.ndarray], sort: bool = True) -> np.ndarray:
    """DC:137-203, full-async branch: one future per host on the shared pool (DC:148-156), join
    in index order (DC:158-159), concatenate in shard order (DC:161-164), sort (DC:195)."""
    futures: List[Future] = [threadPool.submit(sendRequest, channelList[i], featIdsList[i], featWtsList[i])
                             for i in range(len(channelList))]
    responses = [f.result() for f in futures]
    full = np.concatenate(responses) if responses else np.zeros(0, np.float32)
    if sort:
        bits = full.view(np.int32).copy()
        bits[np.isnan(full)] = np.int32(0x7FC00000)
        bits[bits < 0] ^= np.int32(0x7FFFFFFF)
        full = full[np.argsort(bits, kind="stable")]          # Float.compareTo order
    return full


def predict(engine: "binding.Engine", ids: np.ndarray, wts: np.ndarray) -> np.ndarray:
    """The whole of processARequest (minus the sort) as ONE engine call: the engine splits
    the candidates over its GPUs by the DC:46-55 rule and merges in order."""
    return engine.predict(np.ascontiguousarray(ids, np.int64), np.ascontiguousarray(wts, np.float32))


# ---------------------------------------------------------------------------- one process per GPU
def shard_for_rank(n_cand: int, world: int, rank: int):
    """Candidate range of ``rank`` under the DC:46-55 rule (remainder to the last rank)."""
    return binding.partition_bounds(n_cand, world)[rank]


def processARequestDistributed(score_fn, ids, wts, field_num: int, group=None, src: int = 0):
    """processARequest (DC:137-203, minus the sort) when every GPU has its own process
    (``torchrun``; NCCL on the GPU box, gloo in CPU tests): rank ``src`` holds the request, every
    rank scores its DC:46-55 shard with ``score_fn(ids[n,F], wts[n,F]) -> float32[n]`` (on a GPU
    box: ``engine.predict``), and the shards come back to ``src`` in rank order, so output index j
    is candidate j (DC:161-164).  Returns the merged scores on ``src`` and ``None`` elsewhere.
    The only collectives are the scatter of inputs and the gather of scores; the scoring itself
    needs no exchange."""
    import torch
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    meta = [int(np.asarray(ids).size // field_num) if rank == src else 0]
    dist.broadcast_object_list(meta, src=src, group=group)
    n = meta[0]
    bounds = binding.partition_bounds(n, world)
    a, b = bounds[rank]
    my_ids = torch.empty(((b - a) * field_num,), dtype=torch.int64)
    my_wts = torch.empty(((b - a) * field_num,), dtype=torch.float32)
    if rank == src:
        fi = torch.from_numpy(np.ascontiguousarray(ids, np.int64).reshape(-1))
        fw = torch.from_numpy(np.ascontiguousarray(wts, np.float32).reshape(-1))
        parts_i = [fi[s * field_num:e * field_num].contiguous() for s, e in bounds]
        parts_w = [fw[s * field_num:e * field_num].contiguous() for s, e in bounds]
    # shards are ragged (the last one takes the remainder): point-to-point instead of scatter
    if rank == src:
        reqs = []
        for r in range(world):
            if r == src:
                my_ids.copy_(parts_i[r]); my_wts.copy_(parts_w[r])
            else:
                reqs += [dist.isend(parts_i[r], r, group=group), dist.isend(parts_w[r], r, group=group)]
        [q.wait() for q in reqs]
    else:
        dist.recv(my_ids, src, group=group)
        dist.recv(my_wts, src, group=group)
    local = torch.from_numpy(np.ascontiguousarray(
        score_fn(my_ids.numpy().reshape(-1, field_num), my_wts.numpy().reshape(-1, field_num)), np.float32))
    if rank == src:
        out = torch.empty(n, dtype=torch.float32)
        for r, (s, e) in enumerate(bounds):
            if r == src:
                out[s:e] = local
            else:
                dist.recv(out[s:e], r, group=group)
        return out.numpy()
    dist.send(local, src, group=group)
    return None
