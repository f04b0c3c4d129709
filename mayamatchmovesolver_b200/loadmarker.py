"""Marker / camera file readers -> arrays for `mmb_scene_desc`.

The step just before the solve path: the reference's Load Marker tool reads 2D tracks into
MarkerData objects (python/mmSolver/utils/loadmarker/formats/*.py) and creates Maya marker
nodes from them.  Here the same files become `MarkerTrack`s and then markers of a
`SceneBuilder`, without Maya.  Formats (behaviour follows the reference's loaders, quirks
included; tests/golden/loadmarker_golden.json holds what the reference's own loaders return
for its tests/data files):

  .uv   UV Track v1 (text) and v2 / v3 / v4 (JSON)        formats/uvtrack.py:150-588
  .rz2  MatchMover TrackPoints                             formats/rz2.py:37-146
  .txt  3DEqualizer 2D tracks (pixels / image size)        formats/tdetxt.py:77-208
  .2dt / .txt  PFTrack 2D tracks (pixel centres + 0.5)     formats/pftrack2dt.py:97-239
  .mmcamera    camera attribute curves (JSON)              tools/loadmmcamera
"""
import json
import math
import os
import re


class ParserError(ValueError):
    pass


class MarkerTrack:
    """MarkerData (utils/loadmarker/markerdata.py:30-144): per-frame x, y (0..1, origin
    bottom-left), enable, weight; optional bundle position and locks"""

    def __init__(self, name=None):
        self.name = name
        self.group_name = None
        self.id = None
        self.color = None
        self.x, self.y, self.enable, self.weight = {}, {}, {}, {}
        self.bundle = (None, None, None)
        self.bundle_lock = (None, None, None)

    def frames(self):
        return sorted(self.x)

    def as_dict(self):
        key = lambda d: {str(k): d[k] for k in sorted(d)}
        return dict(name=self.name, group_name=self.group_name, id=self.id, color=self.color, x=key(self.x),
                    y=key(self.y), enable=key(self.enable), weight=key(self.weight), bundle=list(self.bundle),
                    bundle_lock=list(self.bundle_lock))


def _file_info(marker_distorted=False, marker_undistorted=False, bundle_positions=False, camera_field_of_view=None):
    # utils/loadmarker/fileinfo.py:27-63
    return dict(marker_distorted=marker_distorted, marker_undistorted=marker_undistorted,
                bundle_positions=bundle_positions, camera_field_of_view=camera_field_of_view or [])


def _fill_occluded(track, frames, weight_quirk):
    """Frames between the first and last tracked frame that have no data are disabled.
    uvtrack.py / pftrack2dt.py test `int(...) is False`, which is never true, so they leave
    the weight alone (weight_quirk=True); tdetxt.py compares a bool and zeroes the weight."""
    have = set(frames)
    for frame in range(min(frames), max(frames) + 1):
        on = frame in have
        track.enable[frame] = int(on)
        if not on and not weight_quirk:
            track.weight[frame] = 0.0


def _strip_comments(lines):
    out = []
    for line in lines:
        line = line.strip()
        if line.startswith("#"):
            continue
        out.append(line.partition("#")[0])
    return out


def _int_or_none(text):
    try:
        return int(text)
    except ValueError:
        return None


def _float_or_none(text):
    try:
        return float(text)
    except ValueError:
        return None


# ---- .uv ------------------------------------------------------------------------------
def uv_format_version(path):
    with open(path) as fh:
        try:
            data = json.load(fh)
        except ValueError:
            data = {}
    if len(data) == 0:
        return 1
    return data.get("version", -1)


def _read_uv_v1(path):
    with open(path) as fh:
        lines = fh.readlines()
    if not lines:
        raise OSError("No contents in the file: %s" % path)
    count = int(lines[0])
    if count < 1:
        raise ParserError("No points exist.")
    tracks = []
    idx = 1
    for _ in range(count):
        track = MarkerTrack(lines[idx].strip())
        idx += 1
        num_frames = int(lines[idx])
        if num_frames <= 0:
            idx += 1
            continue
        frames = []
        remaining = num_frames
        while remaining > 0:
            idx += 1
            line = lines[idx].strip()
            if not line:
                break
            remaining -= 1
            parts = line.split()
            if len(parts) != 4:
                raise ParserError("File invalid, there must be 4 numbers in a line: %r" % line)
            frame = int(parts[0])
            track.weight[frame] = float(parts[3])
            track.x[frame] = float(parts[1])
            track.y[frame] = float(parts[2])
            frames.append(frame)
        _fill_occluded(track, frames, weight_quirk=True)
        tracks.append(track)
        idx += 1
    return _file_info(marker_undistorted=True), tracks


def _read_uv_json(path, undistorted, with_3d_pos):
    pos_key = "pos" if (undistorted is None or undistorted) else "pos_dist"
    with open(path) as fh:
        data = json.load(fh)
    tracks = []
    for point in data.get("points", []):
        track = MarkerTrack(point.get("name"))
        track.group_name = point.get("set_name")
        track.id = point.get("id")
        if with_3d_pos:
            p3 = point.get("3d")
            track.bundle = (p3.get("x"), p3.get("y"), p3.get("z"))
            track.bundle_lock = (p3.get("x_lock"), p3.get("y_lock"), p3.get("z_lock"))
        per_frame = point.get("per_frame", [])
        if not per_frame:
            continue
        frames = []
        for entry in per_frame:
            frame = entry.get("frame")
            u, v = entry.get(pos_key)
            frames.append(frame)
            track.x[frame] = u
            track.y[frame] = v
            track.weight[frame] = entry.get("weight")
            track.enable[frame] = True
        _fill_occluded(track, frames, weight_quirk=True)
        tracks.append(track)
    return data, tracks


def _uv_camera_fov(data):
    # uvtrack.py:363-391: angle of view per frame from focal length and film back
    cam = data.get("camera", {})
    fbx, fby = cam["film_back_cm"]
    out = []
    for entry in cam.get("per_frame", []):
        focal = entry["focal_length_cm"]
        out.append((entry["frame"], math.degrees(2.0 * math.atan(fbx / (2.0 * focal))),
                    math.degrees(2.0 * math.atan(fby / (2.0 * focal)))))
    return out


def read_uv(path, undistorted=None, with_3d_pos=True):
    version = uv_format_version(path)
    if version == 1:
        return _read_uv_v1(path)
    if version == 2:
        _, tracks = _read_uv_json(path, True, False)
        return _file_info(marker_undistorted=True), tracks
    if version == 3:
        _, tracks = _read_uv_json(path, undistorted, with_3d_pos)
        return _file_info(True, True, True), tracks
    if version == 4:
        data, tracks = _read_uv_json(path, undistorted, with_3d_pos)
        return _file_info(True, True, True, _uv_camera_fov(data)), tracks
    raise ParserError("Could not determine format version for UV Track file.")


# ---- .rz2 -----------------------------------------------------------------------------
def read_rz2(path):
    with open(path) as fh:
        text = fh.read()
    idx = text.find("imageSequence")
    if idx == -1:
        raise ParserError("no 'imageSequence' block")
    start = text.find("{", idx + 1)
    end = text.find("}", start + 1)
    if start == -1 or end == -1:
        raise ParserError("unterminated 'imageSequence' block")
    seq = text[start + 1:end].strip()
    parts = seq.split()
    x_res, y_res = int(parts[0]), int(parts[1])
    rng = re.search(r".*b\(\s(\d*)\s(\d*)\s(\d*)\s\)", seq)
    if rng is None:
        raise ParserError("no frame range in %r" % seq)
    first, last, step = (int(g) for g in rng.groups())
    frames = range(first, last + 1, step)
    tracks = []
    idx = end
    while True:
        idx = text.find("pointTrack", idx + 1)
        if idx == -1:
            break
        start = text.find("{", idx + 1)
        if start == -1:
            break
        end = text.find("}", start + 1)
        if end == -1:
            break
        m = re.search(r'pointTrack\s*"(.*)".*', text[idx:start])
        if m is None:
            continue
        track = MarkerTrack(m.group(1))
        track.weight[first] = 1.0
        for frame in frames:
            track.enable[frame] = 0
        for line in text[start + 1:end].splitlines():
            cols = line.split()
            if not cols:
                continue
            frame = int(cols[0])
            # MatchMover's origin is top-left
            track.enable[frame] = int(frame in frames)
            track.x[frame] = float(cols[1]) / x_res
            track.y[frame] = ((float(cols[2]) / y_res) * -1) + 1.0
        tracks.append(track)
    return _file_info(), tracks


# ---- 3DE .txt / PFTrack .2dt ----------------------------------------------------------------
def read_3de_txt(path, image_width=None, image_height=None):
    inv_w = 1.0 / (1.0 if image_width is None else image_width)
    inv_h = 1.0 / (1.0 if image_height is None else image_height)
    with open(path) as fh:
        lines = fh.readlines()
    if not lines:
        raise OSError("No contents in the file: %s" % path)
    lines = _strip_comments(lines)
    count = _int_or_none(lines[0].strip())
    if count is None:
        raise ParserError("Invalid file format.")
    if count < 1:
        raise ParserError("No points exist.")
    tracks = []
    idx = 1
    for _ in range(count):
        track = MarkerTrack(lines[idx].strip())
        idx += 1
        track.color = _int_or_none(lines[idx].strip())
        if track.color is None:
            raise ParserError("Invalid file format.")
        idx += 1
        num_frames = _int_or_none(lines[idx].strip())
        if num_frames is None:
            raise ParserError("Invalid file format.")
        if num_frames <= 0:
            idx += 1
            continue
        frames = []
        remaining = num_frames
        while remaining > 0:
            idx += 1
            line = lines[idx].strip()
            if not line:
                break
            remaining -= 1
            cols = line.split()
            if len(cols) != 3:
                raise ParserError("File invalid, there must be 3 numbers in line: %r" % line)
            frame, px, py = _int_or_none(cols[0]), _float_or_none(cols[1]), _float_or_none(cols[2])
            if frame is None or px is None or py is None:
                raise ParserError("Invalid file format.")
            track.weight[frame] = 1.0
            track.x[frame] = px * inv_w
            track.y[frame] = py * inv_h
            frames.append(frame)
        _fill_occluded(track, frames, weight_quirk=False)
        tracks.append(track)
        idx += 1
    return _file_info(), tracks


def read_pftrack_2dt(path, image_width=None, image_height=None):
    inv_w = 1.0 / (1.0 if image_width is None else image_width)
    inv_h = 1.0 / (1.0 if image_height is None else image_height)
    with open(path) as fh:
        lines = fh.readlines()
    if not lines:
        raise OSError("No contents in the file: %s" % path)
    lines = _strip_comments(lines)
    tracks = []
    i = 0
    while i < len(lines):
        line = lines[i]
        i += 1
        if not (line.startswith('"') and line.endswith('"')):
            continue
        track = MarkerTrack(line[1:-1])
        line = lines[i]
        if _int_or_none(line) is not None or (line.startswith('"') and line.endswith('"')):
            i += 1  # clip number or camera name
        else:
            raise ParserError("File invalid, expecting a camera name (string) in line: %r" % line)
        count = _int_or_none(lines[i])
        if count is None:
            raise ParserError("File invalid, expecting a number of frames (integer) in line: %r" % lines[i])
        i += 1
        frames = []
        for _ in range(count):
            cols = lines[i].split(" ")
            if len(cols) not in (4, 5):
                raise ParserError("File invalid, there must be 4 or 5 numbers in line: %r" % lines[i])
            frame, px, py = _int_or_none(cols[0]), _float_or_none(cols[1]), _float_or_none(cols[2])
            if frame is None or px is None or py is None:
                raise ParserError("Invalid file format.")
            track.weight[frame] = 1.0
            track.x[frame] = (px + 0.5) * inv_w   # pixel centres
            track.y[frame] = (py + 0.5) * inv_h
            frames.append(frame)
            i += 1
        _fill_occluded(track, frames, weight_quirk=True)
        tracks.append(track)
    return _file_info(), tracks


READERS = {".uv": [read_uv], ".rz2": [read_rz2], ".2dt": [read_pftrack_2dt], ".txt": [read_3de_txt, read_pftrack_2dt]}


def read(path, **kwargs):
    """Dispatch on the file extension; a .txt is tried as 3DEqualizer first, then PFTrack
    (utils/loadmarker/fileutils.py:42-80)."""
    ext = os.path.splitext(path)[1].lower()
    if ext not in READERS:
        raise ParserError("unknown marker file extension %r" % ext)
    last = None
    for reader in READERS[ext]:
        try:
            kw = kwargs if reader is not read_rz2 else {}
            if reader is read_uv:
                kw = {k: v for k, v in kwargs.items() if k in ("undistorted", "with_3d_pos")}
            elif reader is not read_rz2:
                kw = {k: v for k, v in kwargs.items() if k in ("image_width", "image_height")}
            return reader(path, **kw)
        except (ParserError, IndexError, ValueError) as e:
            last = e
    raise ParserError("could not read %s: %s" % (path, last))


# ---- .mmcamera ------------------------------------------------------------------------------
MMCAMERA_ATTRS = ("translateX", "translateY", "translateZ", "rotateX", "rotateY", "rotateZ", "focalLength",
                  "filmBackWidth", "filmBackHeight", "filmBackOffsetX", "filmBackOffsetY")


def read_mmcamera(path):
    """-> dict(name, start_frame, end_frame, image, attr={name: {frame: value}}); film back and
    focal length in mm (tools/loadmmcamera reads them the same way)"""
    with open(path) as fh:
        doc = json.load(fh)
    data = doc.get("data", {})
    attrs = {}
    for name, keys in data.get("attr", {}).items():
        attrs[name] = {int(frame): float(value) for frame, value in keys}
    return dict(version=doc.get("version"), name=data.get("name"), start_frame=data.get("start_frame"),
                end_frame=data.get("end_frame"), image=data.get("image", {}), attr=attrs)


# ---- into a scene ---------------------------------------------------------------------------
def add_tracks_to_scene(sb, camera, tracks, bundle_defaults=(0.0, 0.0, -10.0), name_prefix=""):
    """One bundle + marker per track on the frames of `sb`.  Marker positions become
    -0.5..0.5 (the Load Marker tool writes `x - 0.5`, `y - 0.5` onto the marker nodes);
    frames without data are disabled with weight 0.  A track's 3D point (uv v3 / v4) is the
    bundle's start position.  -> [(marker node, bundle node, bundle transform node)]"""
    import numpy as np
    out = []
    for track in tracks:
        x = np.zeros(sb.F)
        y = np.zeros(sb.F)
        w = np.zeros(sb.F)
        en = np.zeros(sb.F, np.uint8)
        for i, frame in enumerate(int(f) for f in sb.frames):
            if frame in track.x and track.enable.get(frame, 1):
                x[i] = track.x[frame] - 0.5
                y[i] = track.y[frame] - 0.5
                wv = track.weight.get(frame)
                w[i] = 1.0 if wv is None else float(wv)
                en[i] = 1
        pos = [bundle_defaults[k] if track.bundle[k] is None else float(track.bundle[k]) for k in range(3)]
        name = name_prefix + (track.name or "track%d" % len(out))
        tfm = sb.add_transform(name=name + "_BND_tfm", tx=pos[0], ty=pos[1], tz=pos[2])
        bnd = sb.add_bundle(tfm, name=name + "_BND")
        mkr = sb.add_marker(camera, bnd, x, y, weight=w, enable=en, name=name + "_MKR")
        out.append((mkr, bnd, tfm))
    return out
