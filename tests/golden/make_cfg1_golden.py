"""Makes the cfg1 fixtures (BASELINE.json configs[0] / SURVEY.md section 8d) from the reference's
own test data; run in the build container (needs /root/reference):

  tests/golden/cfg1_loadmarker_frame41.rz2  frame 41 of the 20 tracks of
      tests/data/match_mover/loadmarker.rz2 that are tracked on that frame (same file format,
      read back by mayamatchmovesolver_b200.loadmarker).  SURVEY.md names frame 0, but only 14
      points are tracked there; frame 41 -- the other end of the shot, also keyed in the
      operahouse test -- has exactly the 20 that BASELINE.json configs[0] asks for.
  tests/golden/cfg1_operahouse.json        camera pose / intrinsics of frame 41
      (tests/test/test_api/test_solve_operahouse.py:62-100) and the bundle positions
      (:166-214) of those tracks
"""
import ast
import json
import os
import re

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
FRAME = 41


def main():
    src = open(os.path.join(REF, "tests/data/match_mover/loadmarker.rz2")).read()
    head, rest = src.split("pointTrack", 1)
    blocks = re.findall(r'pointTrack\s+"([^"]+)"([^{]*)\{([^}]*)\}', "pointTrack" + rest)
    out = [head.rstrip() + "\n"]
    names = []
    for name, colour, body in blocks:
        lines = [l for l in body.strip().splitlines() if l.split() and l.split()[0] == str(FRAME)]
        if not lines or len(names) == 20:
            continue
        names.append(name)
        out.append('pointTrack  "%s" %s\n{\n%s\n}\n' % (name, colour.strip(), lines[0].rstrip()))
    with open(os.path.join(HERE, "cfg1_loadmarker_frame41.rz2"), "w") as fh:
        fh.write("".join(out))
    test = open(os.path.join(REF, "tests/test/test_api/test_solve_operahouse.py")).read()
    m = re.search(r"bnd_positions = (\{.*?\n        \})", test, re.S)
    positions = ast.literal_eval(m.group(1))
    m = re.search(r"cam_data = (\{.*?\n        \})", test, re.S)
    cam = ast.literal_eval(m.group(1))[str(FRAME)]
    data = dict(
        source="tests/data/match_mover/loadmarker.rz2 frame 41; tests/test/test_api/test_solve_operahouse.py:62-100,166-214",
        image_size=[3072, 2304], focal_length_mm=14.0, film_back_mm=[5.4187, 4.0640], rotate_order="zxy",
        frame=FRAME, camera_pose=list(cam), tracks=names,
        bundle_positions={n: list(positions[n.replace(" ", "_")]) for n in names})
    with open(os.path.join(HERE, "cfg1_operahouse.json"), "w") as fh:
        json.dump(data, fh, indent=1)
    print(len(names), "tracks")


if __name__ == "__main__":
    main()
