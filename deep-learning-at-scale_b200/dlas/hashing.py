"""Host-side counter hash (splitmix64), the same function csrc/common.cuh implements on the device.
Used for the per-epoch batch order and for weight initialisation."""
import numpy as np


def splitmix64(x):
    x = np.asarray(x, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = x + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def stream_key(seed, stream):
    with np.errstate(over="ignore"):
        s = np.uint64(seed & 0xFFFFFFFFFFFFFFFF)
        t = np.uint64(stream & 0xFFFFFFFFFFFFFFFF) * np.uint64(0xD1342543DE82EF95)
        return splitmix64(s ^ splitmix64(t))


def uniform(seed, stream, n):
    key = stream_key(seed, stream)
    with np.errstate(over="ignore"):
        h = splitmix64(key + np.arange(n, dtype=np.uint64))
    return (h >> np.uint64(40)).astype(np.float32) * np.float32(2.0 ** -24)


STREAM_PERM = 21


def batch_order(seed, epoch, n_batches):
    """Per-epoch shuffle (Keras fit(shuffle=True), cnn.py:127) at batch granularity: batches stay
    contiguous row blocks (coalesced / TMA-friendly); their order is a hash-keyed permutation."""
    key = stream_key(seed, STREAM_PERM + 1000 * (epoch + 1))
    with np.errstate(over="ignore"):
        keys = splitmix64(key + np.arange(n_batches, dtype=np.uint64))
    return np.argsort(keys, kind="stable").astype(np.int64)
