"""ORACLE — test infrastructure.  The golden-vector case list shared by
``oracle/make_golden.py`` (which runs the real reference) and the tests (which
re-create the same synthetic weights/inputs and compare against the stored vectors)."""

# name -> (model, batch, H, W, input kind, metadata, weight seed, input seed)
CASES = {
    "std_224_b2_none":        ("standard", 2, 224, 224, "unit",   None, 0, 11),
    "std_518_b1_rover":       ("standard", 1, 518, 518, "unit",   {"source": "rover"}, 0, 12),
    "std_256x320_b1_ingalt":  ("standard", 1, 256, 320, "unit",   {"source": "ingenuity", "altitude": 4.0}, 0, 13),
    "std_256_b1_byte_other":  ("standard", 1, 256, 256, "byte",   {"source": "orbiter"}, 0, 14),
    "std_256_b1_signed_none": ("standard", 1, 256, 256, "signed", None, 0, 15),
    "std_160x200_b2_ing":     ("standard", 2, 160, 200, "unit",   {"source": "ingenuity"}, 1, 16),
    "enh_224_b2_none":        ("enhanced", 2, 224, 224, "unit",   None, 0, 21),
    "enh_518_b1_roverh":      ("enhanced", 1, 518, 518, "unit",   {"source": "rover", "height": 1.5}, 0, 22),
    "enh_320x256_b1_sat":     ("enhanced", 1, 320, 256, "unit",   {"source": "satellite"}, 0, 23),
    "enh_1024_b1_rover":      ("enhanced", 1, 1024, 1024, "unit", {"source": "rover"}, 0, 24),
    "std_1024_b1_rover":      ("standard", 1, 1024, 1024, "unit", {"source": "rover"}, 0, 25),
    "cfg4_256_b1_rover":      ("config4",  1, 256, 256, "unit",   {"source": "rover"}, 0, 31),
    "cfg4_1024_b1_rover":     ("config4",  1, 1024, 1024, "unit", {"source": "rover"}, 0, 32),  # BASELINE.json config 4 at its own size
}

# stand-alone refinement.py module cases: name -> (H, W, rgb H, rgb W, seed)
REFINE_CASES = {"refine_96x128": (96, 128, 96, 128, 41), "refine_64x64_rgb128": (64, 64, 128, 128, 42)}
FUSION_CASES = {"fusion_4maps_80x96": (80, 96, 43)}

SUBSAMPLE_TARGET = 48  # stored maps are strided to at most ~48x48 samples


def stride_for(n: int) -> int:
    return max(1, -(-n // SUBSAMPLE_TARGET))


def key_name(k) -> str:
    return f"{k[0]}__{k[1]}"
