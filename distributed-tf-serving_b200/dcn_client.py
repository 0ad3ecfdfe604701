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
