"""Helpers shared by the parity tests."""
import numpy as np

from oracle import restated as R


def keyed(octave, xl, yl):
    return {(int(l), int(x), int(y)): i for i, (l, x, y) in enumerate(zip(octave, xl, yl))}


def level_coords(kps, h, w):
    """Recover integer level coordinates of library keypoints exactly as OpenCV's descriptor stage does
    (cx = rint(pt.x * (1/s_l)))."""
    sizes = R.level_sizes(h, w)
    sc = np.array([sizes[int(l)][2] for l in kps["octave"]], np.float32)
    inv = (np.float32(1.0) / sc).astype(np.float32)
    xl = np.rint((kps["x"] * inv).astype(np.float32)).astype(np.int64)
    yl = np.rint((kps["y"] * inv).astype(np.float32)).astype(np.int64)
    return xl, yl


def compare_orb(kps, desc, ref, h, w):
    """Compare library output with an oracle dict (restated or cv2-packed).  Returns a stats dict."""
    xl, yl = level_coords(kps, h, w)
    ours = keyed(kps["octave"], xl, yl)
    if "xl" in ref:
        rxl, ryl = ref["xl"], ref["yl"]
    else:
        tmp = np.zeros(len(ref["x"]), dtype=[("x", "<f4"), ("y", "<f4"), ("octave", "<i4")])
        tmp["x"], tmp["y"], tmp["octave"] = ref["x"], ref["y"], ref["octave"]
        rxl, ryl = level_coords(tmp, h, w)
    theirs = keyed(ref["octave"], rxl, ryl)
    common = [k for k in theirs if k in ours]
    oi = np.array([ours[k] for k in common], np.int64)
    ri = np.array([theirs[k] for k in common], np.int64)
    st = {
        "n_ours": len(ours), "n_ref": len(theirs),
        "missing": len(set(theirs) - set(ours)), "extra": len(set(ours) - set(theirs)),
        "common": len(common),
    }
    if len(common):
        st["resp_eq"] = int(np.sum(kps["response"][oi] == ref["response"][ri]))
        st["angle_eq"] = int(np.sum(kps["angle"][oi] == ref["angle"][ri]))
        st["angle_maxdiff"] = float(np.max(np.abs(kps["angle"][oi] - ref["angle"][ri])))
        st["pt_eq"] = int(np.sum((kps["x"][oi] == ref["x"][ri]) & (kps["y"][oi] == ref["y"][ri])))
        st["size_eq"] = int(np.sum(kps["size"][oi] == ref["size"][ri]))
        if desc is not None and ref.get("desc") is not None:
            x = desc[oi] ^ ref["desc"][ri]
            st["desc_bits_diff"] = int(np.unpackbits(x).sum())
            st["desc_rows_eq"] = int(np.sum(~x.any(axis=1)))
    return st
