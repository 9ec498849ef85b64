"""Synthetic 8-bit luma sequences and the random-access prediction structure used by tests and bench.

Sequence definition: SURVEY.md 8(d) / BASELINE.md 3 ("motion" and "iid" sets, seed 1234).
Reference lists: POC deltas of the reference's 5-layer random-access RPS for a 16-picture mini-GOP,
restated from the per-GOP-index tables documented at Codec/pd_process.c:1910-2110.
"""
import numpy as np

# gop index (poc % 16) -> (temporal layer, list0 deltas, list1 deltas); delta d means picture poc - d
RA5_RPS = {
    0: (0, [16, 48], [16, 32]),
    8: (1, [8, 16, 24], [-8]),
    4: (2, [4, 8, 12, 20], [-4, -12]),
    12: (2, [4, 8, 12], [-4]),
    2: (3, [2, 4, 10, 18], [-2, -6, -14]),
    6: (3, [2, 4, 6, 14], [-2, -10]),
    10: (3, [2, 4, 10, 18], [-2, -6]),
    14: (3, [2, 4, 6, 14], [-2]),
    1: (4, [1, 9, 8, 17], [-1, -3, -7]),
    3: (4, [1, 3, 2, 11], [-1, -5, -13]),
    5: (4, [1, 5, 4, 13], [-1, -3, -11]),
    7: (4, [1, 3, 6, 7], [-1, -9]),
    9: (4, [1, 9, 8, 17], [-1, -3, -7]),
    11: (4, [1, 3, 2, 11], [-1, -5]),
    13: (4, [1, 5, 4, 13], [-1, -3]),
    15: (4, [1, 3, 6, 7], [-1]),
}
# decode order of gop indices inside a mini-GOP (Codec/pred_structure.c:218-290)
RA5_DECODE_ORDER = [0, 8, 4, 2, 1, 3, 6, 5, 7, 12, 10, 9, 11, 14, 13, 15]


def ref_lists(poc, n_pictures, max_l0, max_l1):
    """(temporal_layer, [list0 pocs], [list1 pocs]) for picture `poc` of an n_pictures sequence.
    References outside the sequence are dropped; an empty list 1 repeats list0[0] (the reference points
    BWD at LAST in that case, pd_process.c:1087). poc 0 is the key frame (no ME)."""
    tl, d0, d1 = RA5_RPS[poc % 16]
    l0 = [poc - d for d in d0 if 0 <= poc - d < n_pictures]
    l1 = [poc - d for d in d1 if 0 <= poc - d < n_pictures]
    l0 = list(dict.fromkeys(l0))[:max_l0]
    l1 = list(dict.fromkeys(l1))[:max_l1]
    if not l0:
        l0 = [max(p for p in range(poc) if p % 16 == 0)] if poc > 0 else []
    if not l1 and l0 and max_l1 > 0:
        l1 = [l0[0]]
    return tl, l0, l1


def decode_order(n_pictures):
    """POCs in decode order, key frame first, mini-GOPs of 16 (pictures past the end are skipped)."""
    order = [0]
    base = 0
    while base + 1 < n_pictures:
        for g in RA5_DECODE_ORDER:
            poc = base + (16 if g == 0 else g)
            if poc < n_pictures:
                order.append(poc)
        base += 16
    return order


def _box5(a):
    """5x5 box filter with edge padding, integer arithmetic (deterministic across numpy builds)."""
    p = np.pad(a.astype(np.uint32), 2, mode="edge")
    c = np.cumsum(np.cumsum(p, axis=0), axis=1)
    c = np.pad(c, ((1, 0), (1, 0)))
    h, w = a.shape
    s = c[5:5 + h, 5:5 + w] - c[0:h, 5:5 + w] - c[5:5 + h, 0:w] + c[0:h, 0:w]
    return ((s + 12) // 25).astype(np.uint8)


class Sequence:
    """Lazy generator of the synthetic luma pictures."""

    RECTS = [(-7, 2), (11, -5), (0, 16), (-24, 0)]  # px / frame

    def __init__(self, width, height, n_pictures, kind="motion", seed=1234, bit_depth=8):
        self.width, self.height, self.n, self.kind, self.seed, self.bit_depth = width, height, n_pictures, kind, seed, bit_depth
        rng = np.random.default_rng(seed)
        if kind == "motion":
            self.field = _box5(rng.integers(0, 256, size=(height + 512, width + 512), dtype=np.uint8))
            # rectangles are 256x192 at 1080p/4K; small test pictures shrink them to a third of the picture
            self.rw, self.rh = min(256, (width // 3) & ~7), min(192, (height // 3) & ~7)
            self.rect_tex = [_box5(rng.integers(0, 256, size=(self.rh, self.rw), dtype=np.uint8)) for _ in self.RECTS]
            self.rect_pos = [(int(rng.integers(0, width - self.rw)), int(rng.integers(0, height - self.rh)))
                             for _ in self.RECTS]

    def luma(self, poc):
        rng = np.random.default_rng([self.seed, poc])
        h, w = self.height, self.width
        if self.kind == "iid":
            if self.bit_depth == 10:
                return (rng.integers(0, 1024, size=(h, w), dtype=np.uint16) >> 2).astype(np.uint8)
            return rng.integers(0, 256, size=(h, w), dtype=np.uint8)
        ox, oy = (3 * poc) % 512, (1 * poc) % 512
        img = self.field[oy:oy + h, ox:ox + w].astype(np.int16)
        for (vx, vy), tex, (px, py) in zip(self.RECTS, self.rect_tex, self.rect_pos):
            x = (px + vx * poc) % (w - self.rw)
            y = (py + vy * poc) % (h - self.rh)
            img[y:y + self.rh, x:x + self.rw] = tex
        img += rng.integers(-2, 3, size=(h, w), dtype=np.int16)
        return np.clip(img, 0, 255).astype(np.uint8)
