"""Programmatic model-graph definitions for the YOLOv7 family.

The reference keeps its graphs as YAML files (cfg/training/*.yaml, consumed by
models/yolo.py:736 parse_model).  This module *generates* the same
``{nc, depth_multiple, width_multiple, anchors, backbone, head}`` dictionaries
from a handful of block builders so that the framework is usable on a machine
that has no copy of the reference tree.  ``tests/test_zoo.py`` checks, when
/root/reference is mounted, that every generated dict equals the parsed YAML.

A row is ``[from, number, module, args]`` exactly as parse_model expects
(SURVEY.md A.3).  ``Model(cfg)`` also accepts a path to any reference YAML.
"""

ANCHORS_P5 = [[12, 16, 19, 36, 40, 28], [36, 75, 76, 55, 72, 146], [142, 110, 192, 243, 459, 401]]
ANCHORS_TINY = [[10, 13, 16, 30, 33, 23], [30, 61, 62, 45, 59, 119], [116, 90, 156, 198, 373, 326]]
ANCHORS_P6 = [[19, 27, 44, 40, 38, 94], [96, 68, 86, 152, 180, 137],
              [140, 301, 303, 264, 238, 542], [436, 615, 739, 380, 925, 792]]


class _Rows:
    """Accumulates rows and remembers absolute indices of named taps."""

    def __init__(self):
        self.rows = []

    def add(self, frm, module, args):
        self.rows.append([frm, 1, module, list(args)])
        return len(self.rows) - 1

    def conv(self, c, k=1, s=1, frm=-1, extra=()):
        return self.add(frm, 'Conv', [c, k, s, *extra])


def _wrap(nc, anchors, rows, n_backbone):
    return {'nc': nc, 'depth_multiple': 1.0, 'width_multiple': 1.0, 'anchors': [list(a) for a in anchors],
            'backbone': rows[:n_backbone], 'head': rows[n_backbone:]}


# ----------------------------------------------------------------------------- yolov7 (P5)
def _elan(g, c, cout, taps, depth=4):
    g.conv(c, 1)
    g.conv(c, 1, frm=-2)
    for _ in range(depth):
        g.conv(c, 3)
    g.add(list(taps), 'Concat', [1])
    return g.conv(cout, 1)


def _elan_head(g, c1, c2, cout, depth=4):
    g.conv(c1, 1)
    g.conv(c1, 1, frm=-2)
    for _ in range(depth):
        g.conv(c2, 3)
    g.add([-i for i in range(1, depth + 3)], 'Concat', [1])
    return g.conv(cout, 1)


def _mp_down(g, c, route=None):
    g.add(-1, 'MP', [])
    g.conv(c, 1)
    g.conv(c, 1, frm=-3)
    g.conv(c, 3, 2)
    g.add([-1, -3] + ([route] if route is not None else []), 'Concat', [1])


def yolov7(nc=80):
    """cfg/training/yolov7.yaml (106 nodes)."""
    g = _Rows()
    g.conv(32, 3, 1)
    g.conv(64, 3, 2)
    g.conv(64, 3, 1)
    g.conv(128, 3, 2)
    t = (-1, -3, -5, -6)
    _elan(g, 64, 256, t)
    _mp_down(g, 128)
    p3 = _elan(g, 128, 512, t)
    _mp_down(g, 256)
    p4 = _elan(g, 256, 1024, t)
    _mp_down(g, 512)
    _elan(g, 256, 1024, t)
    nb = len(g.rows)
    spp = g.add(-1, 'SPPCSPC', [512])
    g.conv(256, 1)
    g.add(-1, 'nn.Upsample', ['None', 2, 'nearest'])
    g.conv(256, 1, frm=p4)
    g.add([-1, -2], 'Concat', [1])
    n63 = _elan_head(g, 256, 128, 256)
    g.conv(128, 1)
    g.add(-1, 'nn.Upsample', ['None', 2, 'nearest'])
    g.conv(128, 1, frm=p3)
    g.add([-1, -2], 'Concat', [1])
    n75 = _elan_head(g, 128, 64, 128)
    _mp_down(g, 128, route=n63)
    n88 = _elan_head(g, 256, 128, 256)
    _mp_down(g, 256, route=spp)
    n101 = _elan_head(g, 512, 256, 512)
    a = g.add(n75, 'RepConv', [256, 3, 1])
    b = g.add(n88, 'RepConv', [512, 3, 1])
    c = g.add(n101, 'RepConv', [1024, 3, 1])
    g.add([a, b, c], 'IDetect', ['nc', 'anchors'])
    return _wrap(nc, ANCHORS_P5, g.rows, nb)


# ----------------------------------------------------------------------------- yolov7-tiny
def yolov7_tiny(nc=80):
    """cfg/training/yolov7-tiny.yaml (78 nodes, LeakyReLU(0.1))."""
    g = _Rows()
    L = ('None', 1, 'nn.LeakyReLU(0.1)')

    def cv(c, k=1, s=1, frm=-1):
        return g.conv(c, k, s, frm=frm, extra=L)

    def block(c, cout):
        cv(c, 1)
        cv(c, 1, frm=-2)
        cv(c, 3)
        cv(c, 3)
        g.add([-1, -2, -3, -4], 'Concat', [1])
        return cv(cout, 1)

    cv(32, 3, 2)
    cv(64, 3, 2)
    block(32, 64)
    g.add(-1, 'MP', [])
    n14 = block(64, 128)
    g.add(-1, 'MP', [])
    n21 = block(128, 256)
    g.add(-1, 'MP', [])
    block(256, 512)
    nb = len(g.rows)
    cv(256, 1)
    cv(256, 1, frm=-2)
    g.add(-1, 'SP', [5])
    g.add(-2, 'SP', [9])
    g.add(-3, 'SP', [13])
    g.add([-1, -2, -3, -4], 'Concat', [1])
    cv(256, 1)
    g.add([-1, -7], 'Concat', [1])
    n37 = cv(256, 1)
    cv(128, 1)
    g.add(-1, 'nn.Upsample', ['None', 2, 'nearest'])
    cv(128, 1, frm=n21)
    g.add([-1, -2], 'Concat', [1])
    n47 = block(64, 128)
    cv(64, 1)
    g.add(-1, 'nn.Upsample', ['None', 2, 'nearest'])
    cv(64, 1, frm=n14)
    g.add([-1, -2], 'Concat', [1])
    n57 = block(32, 64)
    cv(128, 3, 2)
    g.add([-1, n47], 'Concat', [1])
    n65 = block(64, 128)
    cv(256, 3, 2)
    g.add([-1, n37], 'Concat', [1])
    n73 = block(128, 256)
    a = cv(128, 3, 1, frm=n57)
    b = cv(256, 3, 1, frm=n65)
    c = cv(512, 3, 1, frm=n73)
    g.add([a, b, c], 'IDetect', ['nc', 'anchors'])
    return _wrap(nc, ANCHORS_TINY, g.rows, nb)


# ----------------------------------------------------------------------------- yolov7-w6 (P6)
def yolov7_w6(nc=80):
    """cfg/training/yolov7-w6.yaml; nc=83 gives yolov7-w6-custom.yaml (123 nodes, IAuxDetect)."""
    g = _Rows()
    t = (-1, -3, -5, -6)
    g.add(-1, 'ReOrg', [])
    g.conv(64, 3, 1)
    g.conv(128, 3, 2)
    _elan(g, 64, 128, t)
    g.conv(256, 3, 2)
    n19 = _elan(g, 128, 256, t)
    g.conv(512, 3, 2)
    n28 = _elan(g, 256, 512, t)
    g.conv(768, 3, 2)
    n37 = _elan(g, 384, 768, t)
    g.conv(1024, 3, 2)
    _elan(g, 512, 1024, t)
    nb = len(g.rows)
    n47 = g.add(-1, 'SPPCSPC', [512])

    def up(c, route):
        g.conv(c, 1)
        g.add(-1, 'nn.Upsample', ['None', 2, 'nearest'])
        g.conv(c, 1, frm=route)
        g.add([-1, -2], 'Concat', [1])

    up(384, n37)
    n59 = _elan_head(g, 384, 192, 384)
    up(256, n28)
    n71 = _elan_head(g, 256, 128, 256)
    up(128, n19)
    n83 = _elan_head(g, 128, 64, 128)
    g.conv(256, 3, 2)
    g.add([-1, n71], 'Concat', [1])
    n93 = _elan_head(g, 256, 128, 256)
    g.conv(384, 3, 2)
    g.add([-1, n59], 'Concat', [1])
    n103 = _elan_head(g, 384, 192, 384)
    g.conv(512, 3, 2)
    g.add([-1, n47], 'Concat', [1])
    n113 = _elan_head(g, 512, 256, 512)
    outs = [g.conv(c, 3, 1, frm=f) for c, f in ((256, n83), (512, n93), (768, n103), (1024, n113))]
    outs += [g.conv(c, 3, 1, frm=f) for c, f in ((320, n83), (640, n71), (960, n59), (1280, n47))]
    g.add(outs, 'IAuxDetect', ['nc', 'anchors'])
    return _wrap(nc, ANCHORS_P6, g.rows, nb)


def yolov7_w6_custom():
    return yolov7_w6(nc=83)


# ----------------------------------------------------------------------------- yolov7-e6e (P6, E-ELAN)
def _eelan(g, c, cout, head=False):
    """Two parallel ELAN halves reading the same input, joined by Shortcut (E-ELAN)."""
    taps = [-i for i in range(1, 9)] if head else [-1, -3, -5, -7, -8]
    c3 = c // 2 if head else c

    def half(f1, f2):
        g.conv(c, 1, frm=f1)
        g.conv(c, 1, frm=f2)
        for _ in range(6):
            g.conv(c3, 3)
        g.add(list(taps), 'Concat', [1])
        return g.conv(cout, 1)

    half(-1, -2)
    half(-11, -12)
    return g.add([-1, -11], 'Shortcut', [1])


def yolov7_e6e(nc=80):
    """cfg/training/yolov7-e6e.yaml (266 nodes: DownC, Shortcut, 8-way Concat)."""
    g = _Rows()
    g.add(-1, 'ReOrg', [])
    g.conv(80, 3, 1)
    g.add(-1, 'DownC', [160])
    _eelan(g, 64, 160)
    g.add(-1, 'DownC', [320])
    n45 = _eelan(g, 128, 320)
    g.add(-1, 'DownC', [640])
    n67 = _eelan(g, 256, 640)
    g.add(-1, 'DownC', [960])
    n89 = _eelan(g, 384, 960)
    g.add(-1, 'DownC', [1280])
    _eelan(g, 512, 1280)
    nb = len(g.rows)
    n112 = g.add(-1, 'SPPCSPC', [640])

    def up(c, route):
        g.conv(c, 1)
        g.add(-1, 'nn.Upsample', ['None', 2, 'nearest'])
        g.conv(c, 1, frm=route)
        g.add([-1, -2], 'Concat', [1])

    up(480, n89)
    n137 = _eelan(g, 384, 480, head=True)
    n136 = n137 - 1
    up(320, n67)
    n162 = _eelan(g, 256, 320, head=True)
    n161 = n162 - 1
    up(160, n45)
    n187 = _eelan(g, 128, 160, head=True)
    n186 = n187 - 1
    g.add(-1, 'DownC', [320])
    g.add([-1, n162], 'Concat', [1])
    n210 = _eelan(g, 256, 320, head=True)
    g.add(-1, 'DownC', [480])
    g.add([-1, n137], 'Concat', [1])
    n233 = _eelan(g, 384, 480, head=True)
    g.add(-1, 'DownC', [640])
    g.add([-1, n112], 'Concat', [1])
    n256 = _eelan(g, 512, 640, head=True)
    outs = [g.conv(c, 3, 1, frm=f) for c, f in ((320, n187), (640, n210), (960, n233), (1280, n256))]
    outs += [g.conv(c, 3, 1, frm=f) for c, f in ((320, n186), (640, n161), (960, n136), (1280, n112))]
    g.add(outs, 'IAuxDetect', ['nc', 'anchors'])
    return _wrap(nc, ANCHORS_P6, g.rows, nb)


ZOO = {
    'yolov7': yolov7,
    'yolov7-tiny': yolov7_tiny,
    'yolov7-w6': yolov7_w6,
    'yolov7-w6-custom': yolov7_w6_custom,
    'yolov7-e6e': yolov7_e6e,
}


def get_cfg(name, **kw):
    """Return a fresh cfg dict by zoo name (``'yolov7'``, ``'yolov7-tiny'``, ...)."""
    key = name[:-5] if name.endswith('.yaml') else name
    key = key.split('/')[-1]
    if key not in ZOO:
        raise KeyError(f'unknown zoo model {name!r}; known: {sorted(ZOO)}')
    return ZOO[key](**kw)
