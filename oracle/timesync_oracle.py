"""TEST INFRASTRUCTURE ONLY -- restatement of the reference's multi-camera frame synchronisation,
CompositeImage::frame_indices_from_timepairs (/root/reference/source/image.cpp:110-196) and the interval selection
around it (:83-98), in plain Python. The driver's own implementation (host/frame_source.cpp) is written differently
(a per-grid-point search); tests/test_frame_source.py checks on random timelines that both select the same frames.
"""
import math

TIME_EPSILON = 1e-10  # source/include/image.hpp


def c_round(x):
    """std::round: halves away from zero (Python's round() goes to even)."""
    return int(math.floor(abs(x) + 0.5)) * (1 if x >= 0 else -1)


def frame_indices_from_timepairs(timepairs, step, threshold, frames, camera_time, time):
    """timepairs: per camera a list of (time, index) sorted by time; appends to frames / camera_time / time."""
    min_time = min(tp[0][0] for tp in timepairs)            # :120-125
    max_time = max(tp[-1][0] for tp in timepairs)
    if step == 0:                                           # :128-137
        if max_time - min_time < TIME_EPSILON:
            step = 1.0
        else:
            for tp in timepairs:
                min_diff = tp[-1][0] - tp[0][0]
                for a, b in zip(tp, tp[1:]):
                    min_diff = min(b[0] - a[0], min_diff)
                step = max(min_diff, step)
    if threshold == 0:                                      # :139
        threshold = step
    min_time -= step                                        # :142-143
    max_time += step
    nslots = c_round((max_time - min_time) / step) + 1      # :145
    ncam = len(timepairs)
    grid = [[1.01 * threshold, 0] for _ in range(nslots * ncam)]   # :149
    for cam, tp in enumerate(timepairs):                    # :152-165
        for t, index in tp:
            slot = c_round((t - min_time) / step)
            for d in (-1, 0, 1):
                cell = grid[ncam * (slot + d) + cam]
                delta = t - min_time - (slot + d) * step
                if abs(delta) + TIME_EPSILON < abs(cell[0]):
                    cell[0], cell[1] = delta, index
    last_delta = 0.0                                        # :168-195
    for slot in range(1, nslots - 1):
        ftime = min_time + slot * step
        idx, ctime, total = [], [], 0.0
        for cam in range(ncam):
            delta, index = grid[ncam * slot + cam]
            if abs(delta) > threshold + TIME_EPSILON:
                break
            idx.append(index)
            ctime.append(ftime + delta)
            total += abs(delta)
        if len(idx) == ncam:
            if not frames or idx != frames[-1]:
                frames.append(idx)
                camera_time.append(ctime)
                time.append(ftime)
            elif total + TIME_EPSILON < last_delta:
                time[-1] = ftime
            last_delta = total


def composite_frames(timelines, intervals):
    """timelines: per camera the sorted frame times; intervals: (start, stop, step, threshold) tuples (:83-98).
    Returns (frames, camera_time, time)."""
    frames, camera_time, time = [], [], []
    for start, stop, step, threshold in intervals:
        timepairs = [[(t, i) for i, t in enumerate(tl) if start <= t <= stop] for tl in timelines]
        if any(not tp for tp in timepairs):
            continue
        frame_indices_from_timepairs(timepairs, step, threshold, frames, camera_time, time)
    return frames, camera_time, time
