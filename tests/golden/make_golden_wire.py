"""Golden fixture for SURVEY 8(f3): the REFERENCE's serialise_np / deserialise_np (utilities_ros.py:7-45), imported
unmodified from /root/reference with std_msgs / planner_experiments.msg / plan_lib stubbed by field-only classes
(ROS is not installed).  Run in the BUILD container only:

    python tests/golden/make_golden_wire.py   ->  tests/golden/ref_wire_agent_info.npz
"""
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/ROS/src/planner_experiments/src/utilities_ROS/utilities_ros.py"


class _Dim(object):
    def __init__(self):
        self.label, self.size, self.stride = "", 0, 0


class _Layout(object):
    def __init__(self):
        self.dim, self.data_offset = [], 0


class _F32(object):
    def __init__(self):
        self.layout, self.data = _Layout(), []


class _Info(object):
    def __init__(self):
        self.header, self.data = None, []


def load_reference():
    for name, attrs in (("std_msgs", {}), ("std_msgs.msg", dict(Float32MultiArray=_F32, MultiArrayDimension=_Dim)),
                        ("planner_experiments", {}), ("planner_experiments.msg", dict(agent_info=_Info)),
                        ("plan_lib", {}), ("plan_lib.utilities", dict(load_var=lambda p: None))):
        m = types.ModuleType(name)
        for k, v in attrs.items():
            setattr(m, k, v)
        sys.modules[name] = m
    spec = importlib.util.spec_from_file_location("ref_utilities_ros", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    ref = load_reference()
    rng = np.random.default_rng(5)
    fields = [rng.normal(size=(21, 9)), rng.normal(size=(20, 2)), rng.normal(size=(4, 3, 2)), rng.normal(size=(7,))]
    msg = ref.serialise_np(fields)
    out = {"n_fields": np.array(len(msg.data))}
    for i, f in enumerate(msg.data):
        out["in%d" % i] = fields[i]
        out["data%d" % i] = np.asarray(f.data)
        out["offset%d" % i] = np.array(f.layout.data_offset)
        out["sizes%d" % i] = np.array([d.size for d in f.layout.dim])
        out["strides%d" % i] = np.array([d.stride for d in f.layout.dim])
        out["labels%d" % i] = np.array([d.label for d in f.layout.dim])
    back = ref.deserialise_np(msg)
    for i, b in enumerate(back):
        out["back%d" % i] = b
    np.savez_compressed(os.path.join(HERE, "ref_wire_agent_info.npz"), **out)
    print("fields", len(msg.data), [b.shape for b in back])


if __name__ == "__main__":
    main()
