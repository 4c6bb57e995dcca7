"""CPU: SURVEY 8(f3) wire-format adaptor against the reference's own serialise_np / deserialise_np
(tests/golden/ref_wire_agent_info.npz, made by tests/golden/make_golden_wire.py from the unmodified reference with
field-only stand-ins for the ROS message classes)."""
import numpy as np

from conftest import golden
from colaborativempc_b200 import wire


def test_layout_matches_reference_serialise_np():
    g = golden("ref_wire_agent_info.npz")
    n = int(g["n_fields"])
    fields = [g["in%d" % i] for i in range(n)]
    msg = wire.serialise_np(fields)
    assert len(msg.data) == n
    for i, f in enumerate(msg.data):
        assert f.layout.data_offset == int(g["offset%d" % i]) == 0
        assert [d.size for d in f.layout.dim] == list(g["sizes%d" % i])
        assert [d.stride for d in f.layout.dim] == list(g["strides%d" % i])      # prod(shape[k:]), utilities_ros.py:29
        assert [d.label for d in f.layout.dim] == list(g["labels%d" % i])
        # the reference hands float64 to a Float32MultiArray; ROS packs float32: same values after that cast
        assert f.data.dtype == np.float32
        assert np.array_equal(f.data, g["data%d" % i].astype(np.float32))
    back = wire.deserialise_np(msg)
    for i, b in enumerate(back):
        assert b.shape == g["back%d" % i].shape
        assert np.array_equal(b, g["back%d" % i].astype(np.float32))


def test_publish_receive_round_trip():
    """LPV_ROS_main.py:71-77: publish xPred, the callback keeps columns (7, 8); float32 on the wire."""
    rng = np.random.default_rng(0)
    N, B = 20, 5
    xpred = rng.normal(size=(B, N + 1, 9)) * 10
    msgs = wire.publish_states(xpred)
    assert len(msgs) == B and all(len(m.data) == 1 for m in msgs)
    nbr = wire.receive_states(msgs[1:4], N)
    assert nbr.shape == (N + 1, 3, 2) and nbr.dtype == np.float64
    assert np.array_equal(nbr, np.swapaxes(xpred[1:4][:, :, [7, 8]], 0, 1).astype(np.float32).astype(np.float64))
    assert np.abs(nbr - np.swapaxes(xpred[1:4][:, :, [7, 8]], 0, 1)).max() < 4e-6 * 10 * 5


def test_foreign_message_classes_are_used():
    class Dim(object):
        def __init__(self):
            self.label, self.size, self.stride = "", 0, 0

    msg = wire.serialise_np([np.zeros((2, 3))], classes=dict(MultiArrayDimension=Dim))
    assert isinstance(msg.data[0].layout.dim[0], Dim) and msg.data[0].layout.dim[1].stride == 3
