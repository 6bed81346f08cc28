"""CPU restatement of the K7 tile aggregate (TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's
CPU leg may import this).

The reference has no arithmetic for this stage: it would run an aggregation as one more task chained with
`AddPipelineTask` (/root/reference/APIs/1.0/Common/task_management/distributed_api_task.py:67-100) whose body is the
API author's own code.  **Parity unpinned by the reference**; the definition below is this repo's and
the CUDA kernel must equal it bit for bit - every output is an integer or a copied float.

    labels[i], label_prob[i] = top-1 class / probability of chip i
    hist[k]                  = number of chips whose top-1 class is k
    mass_q32[k]              = sum over chips and their five entries with class k of round(p * 2**32), p clamped to [0, 1]
"""
import numpy as np


def tile_summary(top_idx5: np.ndarray, top_prob5: np.ndarray, num_classes: int):
    idx = np.asarray(top_idx5, np.int64).reshape(-1, 5)
    prob = np.asarray(top_prob5, np.float32).reshape(-1, 5)
    labels = idx[:, 0].astype(np.int32)
    label_prob = prob[:, 0].copy()
    hist = np.zeros(num_classes, np.uint32)
    ok0 = (idx[:, 0] >= 0) & (idx[:, 0] < num_classes)
    np.add.at(hist, idx[ok0, 0], 1)
    p = np.clip(np.nan_to_num(prob, nan=0.0), 0.0, 1.0).astype(np.float32)
    # p * 2^32 is exact in float32 (a power-of-two scale); its value is an integer below 2^32 + 1
    q = np.rint(p.astype(np.float64) * 4294967296.0).astype(np.uint64)
    mass = np.zeros(num_classes, np.uint64)
    ok = (idx >= 0) & (idx < num_classes)
    np.add.at(mass, idx[ok], q[ok])
    return labels, label_prob, hist, mass
