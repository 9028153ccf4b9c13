"""Stand-in for `matplotlib` (absent; plots are not part of the tracked path).  dpvo/dpvo.py:5,10 and dpvo/net.py:24 import it at
module top; only the visualisation helpers (draw_img_matching_*, save_inlier_ratio_record's PNG) call into it."""


def use(*a, **k):
    pass
