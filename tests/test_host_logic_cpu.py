"""CPU: host-side logic added around the training step (no kernels): CSR offsets of a ragged object batch, the NUMA binding
helper, the stream bookkeeping of the weight-gradient queues."""
import os

import torch

from objcavit_b200.engine import ObjectBatch
from objcavit_b200.sharding import bind_host_to_gpu


def test_object_batch_offsets_are_csr_and_uniform_offsets_cover_the_padded_layout():
    rows = torch.zeros(2 + 5 + 1, 4)
    ob = ObjectBatch(rows, [2, 5, 1])
    assert ob.offsets.dtype == torch.int32 and ob.offsets.tolist() == [0, 2, 7, 8] and ob.offsets_host == [0, 2, 7, 8]
    assert ob.n_max == 5 and ob.uniform_offsets(5).tolist() == [0, 5, 10, 15]
    again = ObjectBatch(rows, [2, 5, 1])                       # same counts: same values (on a GPU: the cached tensor)
    assert again.offsets.tolist() == ob.offsets.tolist()


def test_bind_host_to_gpu_never_raises_and_keeps_a_valid_affinity():
    before = os.sched_getaffinity(0)
    try:
        got = bind_host_to_gpu(0)                                  # no NVML / no GPU here: must report "nothing changed"
        assert got is None or (len(got) > 0 and set(got) <= before)
        assert len(os.sched_getaffinity(0)) > 0
    finally:
        os.sched_setaffinity(0, before)


def test_join_side_waits_only_for_queues_used_since_the_last_join():
    import objcavit_b200.engine_train as et
    assert isinstance(et._side_dirty, set)
    et._side_dirty.clear()
    et.join_side(torch.device("cpu")) if not et.ASYNC_WGRAD else None      # nothing dirty: nothing to wait for, no CUDA call
    assert not et._side_dirty


def test_param_views_share_the_version_counter():
    """the kernels' parameter views must move their ``_version`` with in-place updates of the Parameter."""
    import torch
    from objcavit_b200.config import builtin_args
    from objcavit_b200.modules import build_head
    head = build_head(builtin_args("nyu_adabins"))
    k = "conv_out.0.bias"
    v0 = head._params()[k]._version
    with torch.no_grad():
        head.conv_out[0].bias.add_(1.0)
    assert head._params()[k]._version > v0
    sd = {n: t.clone() for n, t in head.state_dict().items()}
    v1 = head._params()[k]._version
    head.load_state_dict(sd)
    assert head._params()[k]._version > v1


def test_object_frontend_all_none_batch():
    """Yolov7Wrapper returns None for an image without detections; a whole batch of them (batch size 1 at validation) must give
    the placeholders, not raise."""
    from objcavit_b200.evaluation import object_frontend
    assert object_frontend([None, None]) == ([None, None], [None, None])
    assert object_frontend([None]) == ([None], [None])
