"""A batch of registrations as one flat file, the input of examples/batch_pipeline.cpp (the C++ caller of the batch entry
points).  Little-endian:
    8 bytes  "NDTBJOB1"
    int32    n_targets, n_sources, n_items, stride_bytes (12: packed x y z, 16: x y z w)
    int64    target_n[n_targets], source_n[n_sources]
    int32    item_target[n_items], item_source[n_items]
    float32  guesses[n_items][16]        (column-major 4x4, as ndtb_batch_run takes them)
    clouds   every target, then every source, records of stride_bytes
"""
import numpy as np

MAGIC = b"NDTBJOB1"


def _records(cloud, stride):
    c = np.ascontiguousarray(np.asarray(cloud, dtype=np.float32))
    if stride == 12:
        return np.ascontiguousarray(c[:, :3])
    if c.shape[1] == 3:
        c = np.concatenate([c, np.zeros((c.shape[0], 1), np.float32)], axis=1)
    return np.ascontiguousarray(c[:, :4])


def write_job(path, targets, sources, item_target, item_source, guesses, stride=16):
    assert stride in (12, 16)
    it = np.ascontiguousarray(item_target, dtype=np.int32)
    isrc = np.ascontiguousarray(item_source, dtype=np.int32)
    # guesses arrive as numpy 4x4 matrices (row-major); the file holds what ndtb_batch_run takes: column-major
    G = np.ascontiguousarray(np.asarray(guesses, dtype=np.float32).reshape(len(it), 4, 4).transpose(0, 2, 1)).reshape(len(it), 16)
    with open(path, "wb") as f:
        f.write(MAGIC)
        np.array([len(targets), len(sources), len(it), stride], np.int32).tofile(f)
        np.array([len(t) for t in targets], np.int64).tofile(f)
        np.array([len(s) for s in sources], np.int64).tofile(f)
        it.tofile(f)
        isrc.tofile(f)
        G.tofile(f)
        for c in list(targets) + list(sources):
            _records(c, stride).tofile(f)
    return path


def read_results(path):
    """The result file batch_pipeline writes with --out: n_items x (16 T, iterations, converged, score, trans_probability, n_evals)
    as float64 rows."""
    return np.fromfile(path, dtype=np.float64).reshape(-1, 21)
