"""Piecewise-arc track table (host side, start-up only).

The hot path reads a track only through the per-segment table ``PointAndTangent[:, :, lane]`` =
(x_end, y_end, psi_end, s_start, length, curvature) and ``halfWidth`` (reference:
plan_lib/utilities/misc.py:84-126).  ``TrackMap`` builds that table for the reference's track
specifications (segment length / signed radius pairs, plan_lib/mapManager/track_initialization.py:24-131)
with a plain forward integration of the centre line, and offers the curvilinear -> Cartesian map needed
by the start-up prediction (track_initialization.py:325-399).  Any object with the same three attributes
(e.g. the reference's own ``Map``) can be handed to the planner instead.
"""
import math

import numpy as np

_PI = math.pi

# (length, signed radius) per segment; radius 0 = straight.  Data restated from
# track_initialization.py:24-131 (the numbers define the tracks, they are not code).
_SPECS = {
    "Highway": dict(scale=2.0, half_width=0.75, open=True, segs=[
        (0.0, 0.0), (1.0, 0.0), (4.5, 4.5 / (0.5 * _PI)), (2.0, 0.0), (2.5, -2.5 / (0.5 * _PI)), (2.0, 0.0),
        (4.5, 4.5 / _PI), (2.0, 0.0), (5.0, 0.0), (0.0, 0.0)]),
    "oval": dict(scale=1.0, half_width=0.55, open=False, segs=[
        (2.0, 0.0), (5.85, 5.85 / _PI), (4.0, 0.0), (5.85, 5.85 / _PI), (2.0, 0.0)]),
    "oval_mt": dict(scale=1.0, half_width=0.5, open=False, segs=[
        (1.0, 0.0), (4.5, 4.5 / _PI), (2.0, 0.0), (4.5, 4.5 / _PI), (1.0, 0.0)]),
    "TestOpenMap": dict(scale=2.0, half_width=0.5, open=True, segs=[
        (0.0, 0.0), (1.0, 0.0), (4.5, 4.5 / _PI), (2.0, 0.0), (0.0, 0.0)]),
}


def _wrap(a):
    if a < -_PI:
        return 2 * _PI + a
    if a > _PI:
        return a - 2 * _PI
    return a


class TrackMap(object):
    """Drop-in for the attributes of the reference ``Map`` that the LPV path touches."""

    def __init__(self, selectedTrack="Highway", segs=None, half_width=None, open_track=True, scale=1.0):
        if segs is None:
            spec = _SPECS[selectedTrack]
            segs, half_width, open_track, scale = spec["segs"], spec["half_width"], spec["open"], spec["scale"]
        self.lane = 0
        self.open = bool(open_track)
        self.slack = 0.15
        rows = []
        x, y, psi, s = 0.0, 2.0 * half_width, 0.0, 0.0  # tracks start at (0, 2*halfWidth) heading +x
        for (length, radius) in segs:
            length, radius = scale * length, scale * radius
            if radius == 0.0:
                x, y = x + length * math.cos(psi), y + length * math.sin(psi)
                curv = 0.0
            else:
                d = 1.0 if radius >= 0 else -1.0
                r = abs(radius)
                cx, cy = x + r * math.cos(psi + d * _PI / 2), y + r * math.sin(psi + d * _PI / 2)
                span = length / r
                a0 = psi - d * _PI / 2  # direction centre -> start point
                x, y = cx + r * math.cos(a0 + d * span), cy + r * math.sin(a0 + d * span)
                psi = _wrap(psi + d * span)
                curv = 1.0 / radius
            rows.append([x, y, psi, s, length, curv])
            s += length
        if not self.open:  # closing straight back to the start point
            rows.append([0.0, 2.0 * half_width, 0.0, s, math.hypot(rows[-1][0], rows[-1][1] - 2.0 * half_width), 0.0])
        tab = np.asarray(rows, dtype=np.float64)
        self.PointAndTangent = tab[:, :, None].copy()
        self.halfWidth = half_width * np.ones(max(len(segs), tab.shape[0]))
        self.TrackLength = np.array([tab[-1, 3] + tab[-1, 4]])

    # -- lookups with the semantics of misc.py:78-126 (used on the host for start-up data only) --
    def segment(self, s):
        tab = self.PointAndTangent[:, :, self.lane]
        L = tab[-1, 3] + tab[-1, 4]
        while s > L:
            s -= L
        if s < 0:
            s = 0.0
        idx = np.nonzero((s >= tab[:, 3]) & (s < tab[:, 3] + tab[:, 4]))[0]
        if idx.shape[0] != 1:
            raise ValueError("s=%r lies in no unique track segment" % (s,))
        return int(idx[0])

    def getGlobalPosition(self, s, ey, lane=None):
        """(s, e_y) -> (x, y, theta); semantics of track_initialization.py:325-399 for scalar e_y."""
        tab = self.PointAndTangent[:, :, self.lane if lane is None else lane]
        s = float(np.asarray(s).ravel()[0])
        ey = float(np.asarray(ey).ravel()[0])
        L = tab[-1, 3] + tab[-1, 4]
        if self.open:
            if s >= L:
                s -= L
        else:
            while s >= L:
                s -= L
        if s < 0:
            s = 0.0
        i = self.segment(s)
        xs, ys, psi0 = (tab[i - 1, 0], tab[i - 1, 1], tab[i - 1, 2])  # i-1 == -1 wraps like numpy does
        ds = s - tab[i, 3]
        if tab[i, 5] == 0.0:
            psi = tab[i, 2]
            w = ds / tab[i, 4]
            x = (1 - w) * xs + w * tab[i, 0] + ey * math.cos(psi + _PI / 2)
            y = (1 - w) * ys + w * tab[i, 1] + ey * math.sin(psi + _PI / 2)
            return x, y, psi
        r = 1.0 / tab[i, 5]
        d = 1.0 if r >= 0 else -1.0
        ra = abs(r)
        cx, cy = xs + ra * math.cos(psi0 + d * _PI / 2), ys + ra * math.sin(psi0 + d * _PI / 2)
        span = ds / ra
        an = _wrap(d * _PI / 2 + psi0)
        a0 = -(_PI - abs(an)) * (1.0 if an >= 0 else -1.0)
        x = cx + (ra - d * ey) * math.cos(a0 + d * span)
        y = cy + (ra - d * ey) * math.sin(a0 + d * span)
        return x, y, psi0 + d * span
