"""Mirror of the one ``layers.py`` function on the hot path (reference ``layers.py:16-25``)."""


def disp_to_depth(disp, min_depth, max_depth):
    """Sigmoid disparity -> metric depth, ``1 / (1/max + (1/min - 1/max) * disp)``.

    Host-side helper for callers; the forward path fuses the same arithmetic into
    the output kernel (``csrc/pointwise.cu: resize_maps_kernel``)."""
    lo, hi = 1 / max_depth, 1 / min_depth
    scaled = lo + (hi - lo) * disp
    return scaled, 1 / scaled
