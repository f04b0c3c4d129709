"""Generates tests/golden/loadmarker_golden.json with the REFERENCE's own marker-file loaders
(python/mmSolver/utils/loadmarker/formats/*.py, Maya stubbed: they are pure parsing) run on
the reference's own tests/data files.  tests/test_loadmarker_cpu.py checks
mayamatchmovesolver_b200.loadmarker against it.  Build container only:

    python tests/golden/make_loadmarker_golden.py
"""
import json
import os
import sys
import types

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))

FILES = [
    ("uvtrack/test_v1.uv", {}), ("uvtrack/test_v3.uv", {}), ("uvtrack/test_v3_with_3d_point.uv", {}),
    ("uvtrack/test_v4.uv", {}), ("uvtrack/test_v4.uv", {"undistorted": False}), ("uvtrack/eye_fmt1_v001.uv", {}),
    ("uvtrack/eye_fmt2_v001.uv", {}), ("uvtrack/loadmarker_corners.uv", {}), ("uvtrack/stA_with_emptyMarker.uv", {}),
    ("uvtrack/hcw_painting.uv", {}),
    ("match_mover/loadmarker_corners.rz2", {}), ("match_mover/loadmarker.rz2", {}),
    ("3de_v4/loadmarker_corners.txt", {"image_width": 1920.0, "image_height": 1080.0}),
    ("pftrack/pftrack_corners_v001.txt", {"image_width": 1920.0, "image_height": 1080.0}),
    ("pftrack/pftrack_corners_v002.txt", {}),
    ("pftrack/pftrack5_garage.2dt", {"image_width": 2048.0, "image_height": 1152.0}),
]


def main():
    for name in ("maya", "maya.cmds", "maya.mel", "maya.OpenMaya", "maya.api", "maya.api.OpenMaya", "maya.utils"):
        sys.modules[name] = types.ModuleType(name)
    sys.modules["maya"].cmds = sys.modules["maya.cmds"]
    sys.path.insert(0, os.path.join(REF, "python"))
    import mmSolver.utils.loadmarker.formats.pftrack2dt as pf
    import mmSolver.utils.loadmarker.formats.rz2 as rz2
    import mmSolver.utils.loadmarker.formats.tdetxt as tde
    import mmSolver.utils.loadmarker.formats.uvtrack as uv
    loaders = {".uv": uv.LoaderUVTrack, ".rz2": rz2.LoaderRZ2, ".2dt": pf.LoaderPFTrack2DT}

    def kf(k):
        raw = k.get_raw_data()
        return {str(int(f)): raw[f] for f in sorted(raw, key=lambda s: int(s))}

    out = []
    for rel, kwargs in FILES:
        path = os.path.join(REF, "tests", "data", rel)
        ext = os.path.splitext(rel)[1]
        if ext == ".txt":
            loader = tde.Loader3DETXT if rel.startswith("3de") else pf.LoaderPFTrack2DT
        else:
            loader = loaders[ext]
        info, mkrs = loader().parse(path, **kwargs)
        tracks = []
        for m in mkrs:
            tracks.append(dict(name=m.get_name(), group_name=m.get_group_name(), id=m.get_id(), color=m.get_color(),
                               x=kf(m.get_x()), y=kf(m.get_y()), enable=kf(m.get_enable()), weight=kf(m.get_weight()),
                               bundle=[m.get_bundle_x(), m.get_bundle_y(), m.get_bundle_z()],
                               bundle_lock=[m.get_bundle_lock_x(), m.get_bundle_lock_y(), m.get_bundle_lock_z()]))
        # keep the fixture small: long files store the first 3 tracks and per-track key counts
        full = len(json.dumps(tracks)) < 60000
        entry = dict(file=rel, kwargs=kwargs, loader=loader.__name__, info=dict(info._asdict()), num_tracks=len(tracks),
                     names=[t["name"] for t in tracks], key_counts=[len(t["x"]) for t in tracks],
                     sums=[[sum(t["x"].values()), sum(t["y"].values()), sum(float(v or 0) for v in t["weight"].values()),
                            sum(int(v) for v in t["enable"].values())] for t in tracks],
                     tracks=tracks if full else tracks[:3], complete=full)
        entry["info"]["camera_field_of_view"] = [list(v) for v in entry["info"]["camera_field_of_view"]][:5]
        out.append(entry)
        print(rel, kwargs, len(tracks), "tracks", "full" if full else "first 3")
    path = os.path.join(HERE, "loadmarker_golden.json")
    with open(path, "w") as fh:
        json.dump(out, fh, sort_keys=True)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
