"""Wire-format adaptor for the reference's ROS message `agent_info`.

Replaces ROS/src/planner_experiments/src/utilities_ROS/utilities_ros.py:7-45 (serialise_np / deserialise_np) and the
publish/callback pair of LPV_ROS_main.py:71-77 for a node that feeds / consumes the batched planner.  No ROS import:
the message classes below carry exactly the fields of std_msgs/Float32MultiArray and msg/agent_info.msg
("Header header; std_msgs/Float32MultiArray[] data"), so a rospy node copies them field by field, or passes its own
message classes through `classes=`.

Layout written by the reference for every array `field` (utilities_ros.py:13-31):
    data        = field.flatten()  (row-major; float32 ON THE WIRE - the message type is Float32MultiArray)
    data_offset = 0
    dim[k].label = "dim k",  dim[k].size = field.shape[k],  dim[k].stride = prod(field.shape[k:])
"""
import numpy as np


class MultiArrayDimension(object):
    __slots__ = ("label", "size", "stride")

    def __init__(self, label="", size=0, stride=0):
        self.label, self.size, self.stride = label, size, stride


class MultiArrayLayout(object):
    __slots__ = ("dim", "data_offset")

    def __init__(self):
        self.dim, self.data_offset = [], 0


class Float32MultiArray(object):
    __slots__ = ("layout", "data")

    def __init__(self):
        self.layout, self.data = MultiArrayLayout(), []


class AgentInfo(object):
    """msg/agent_info.msg"""
    __slots__ = ("header", "data")

    def __init__(self):
        self.header, self.data = None, []


_CLASSES = dict(agent_info=AgentInfo, Float32MultiArray=Float32MultiArray, MultiArrayDimension=MultiArrayDimension)


def serialise_np(data, header=None, classes=None):
    """utilities_ros.py:7-33.  `data`: list of ndarrays.  The payload is cast to float32, as ROS does when it packs a
    Float32MultiArray (the reference hands it float64 and loses the low bits at publish time)."""
    cls = dict(_CLASSES)
    cls.update(classes or {})
    msg = cls["agent_info"]()
    for field in data:
        field = np.asarray(field)
        proxy = cls["Float32MultiArray"]()
        proxy.layout.data_offset = 0
        proxy.data = field.astype(np.float32).flatten()
        dims = field.shape
        proxy.layout.dim = [cls["MultiArrayDimension"]() for _ in range(field.ndim)]
        for k in range(field.ndim):
            proxy.layout.dim[k].label = "dim " + str(k)
            proxy.layout.dim[k].size = int(dims[k])
            proxy.layout.dim[k].stride = int(np.prod(dims[k:]))
        msg.data.append(proxy)
    if header is not None:
        msg.header = header
    return msg


def deserialise_np(msg):
    """utilities_ros.py:35-45: reshape every field's raw data with the sizes of its layout."""
    return [np.asarray(f.data).reshape(tuple(d.size for d in f.layout.dim)) for f in msg.data]


def publish_states(xpred, classes=None):
    """One `agent_info` per local agent from the batched planner's predictions, as LPV_ROS_main.py:75-77 publishes them
    (`serialise_np([Controller.xPred])`).  xpred: [B, N+1, 9] (host array or a device tensor)."""
    if hasattr(xpred, "detach"):
        xpred = xpred.detach().cpu().numpy()
    return [serialise_np([xpred[b]], classes=classes) for b in range(xpred.shape[0])]


def receive_states(msgs, N):
    """The callback of LPV_ROS_main.py:71-73 for a list of neighbour messages: [N+1, len(msgs), 2] float64 block of
    predicted (x, y) - what `gather_neighbours` / `solve_batch(nbr_xy=...)` consume."""
    out = np.zeros((N + 1, len(msgs), 2))
    for j, m in enumerate(msgs):
        out[:, j, :] = deserialise_np(m)[0][:, [7, 8]]
    return out
