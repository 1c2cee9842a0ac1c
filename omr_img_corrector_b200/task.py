"""Host mirror of the Tauri batch path (packages/app/src-tauri/src/task.rs:19-68 and
thread_pool.rs): the reference fans `add_task` calls out to one OS thread per card, each
emitting `start_running_task`, running omr::correct_default and emitting `task_completed`
with result "finished" | "debatable" | "error".  Here the queued tasks of one parameter set
run through `oics_correct_batch` (upload / estimate + deskew / download pipeline on the GPU)
and the same events are emitted, in task order, with the same payload fields.

Two forms.  `run_tasks`: a task carries the decoded BGR card and `write(output_path, image)` is the
caller's encoder (None = do not write).  `run_file_tasks`: a task carries `input_path` /
`output_path` exactly as the reference's `add_task` does (task.rs:19-30); the JPEG file is read
from disk as bytes, decoded, deskewed and re-encoded on the GPU (`oics_correct_batch_files`:
imread, omr.rs:349, and imwrite(JPEG, 100), omr.rs:447-448, included) and the output bytes are
written to `output_path`.  Files the CUDA decoder does not cover (progressive, Exif-turned,
other formats) come back as "error" unless the caller passes `fallback_decode` (e.g. cv2.imdecode),
in which case they take the pixel path.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional

import numpy as np

from . import api


@dataclass
class Task:
    """task.rs:19-30: one `add_task` call (the parameters are per call in the reference; tasks
    that share them are batched together here)."""
    task_id: int
    card: np.ndarray          # decoded BGR card [h, w, 3]
    output_path: str = ""


def run_tasks(ctx: "api.Context", tasks: List[Task], projection_max_angle: int, projection_angle_step: float,
              projection_max_width: int, projection_max_height: int, hough_min_line_length: float,
              hough_max_line_gap: float, emit: Callable[[str, dict], None],
              write: Optional[Callable[[str, np.ndarray], None]] = None) -> None:
    """Runs the tasks and emits the reference's events through `emit(event_name, payload)`."""
    by_shape = {}
    for t in tasks:
        by_shape.setdefault(tuple(t.card.shape), []).append(t)
    for shape, group in by_shape.items():
        for t in group:
            emit("start_running_task", {"task_id": t.task_id})
        if len(shape) != 3 or shape[2] != 3:
            for t in group:
                emit("task_completed", {"task_id": t.task_id, "output_path": t.output_path, "result": "error"})
            continue
        h, w = shape[:2]
        side = int(np.ceil((w + h) * 0.70710678118654757)) + 2          # CONTAIN bound of any rotation
        outs = [np.empty(side * side * 3, np.uint8) for _ in group] if write else None
        try:
            ang, need, stat, sizes = ctx.correct_batch(np.stack([t.card for t in group]), projection_max_angle,
                                                       projection_angle_step, projection_max_width,
                                                       projection_max_height, hough_min_line_length,
                                                       hough_max_line_gap, outs=outs)
        except api.OicsError:
            for t in group:
                emit("task_completed", {"task_id": t.task_id, "output_path": t.output_path, "result": "error"})
            continue
        for i, t in enumerate(group):
            if stat[i] != 0:
                result = "error"                                            # task.rs:58-62: any Err
            else:
                result = "debatable" if need[i] else "finished"             # task.rs:47-57
                if write:
                    dw, dh = int(sizes[i][0]), int(sizes[i][1])
                    write(t.output_path, outs[i][:dw * dh * 3].reshape(dh, dw, 3))
            emit("task_completed", {"task_id": t.task_id, "output_path": t.output_path, "result": result})


@dataclass
class FileTask:
    """task.rs:19-30 as the reference passes it: paths in, paths out."""
    task_id: int
    input_path: str
    output_path: str


def run_file_tasks(ctx: "api.Context", tasks: List[FileTask], projection_max_angle: int, projection_angle_step: float,
                   projection_max_width: int, projection_max_height: int, hough_min_line_length: float,
                   hough_max_line_gap: float, emit: Callable[[str, dict], None],
                   fallback_decode: Optional[Callable[[bytes], Optional[np.ndarray]]] = None, jpeg_quality: int = 100) -> None:
    """Paths in, paths out, codecs on the device; events as `run_tasks`."""
    params = (projection_max_angle, projection_angle_step, projection_max_width, projection_max_height,
              hough_min_line_length, hough_max_line_gap)
    loaded, by_size = {}, {}
    for t in tasks:
        emit("start_running_task", {"task_id": t.task_id})
        data, w, h = None, 0, 0
        try:
            data = np.fromfile(t.input_path, np.uint8)
            w, h, _ = ctx.jpeg_info(data)
        except OSError:
            data = None
        except api.OicsError:
            w = h = 0                 # readable, but not for the device decoder
        loaded[t.task_id] = data
        by_size.setdefault((w, h), []).append(t)
    leftovers = []
    for (w, h), group in by_size.items():
        if w == 0:
            leftovers += group
            continue
        side = int(np.ceil((w + h) * 0.70710678118654757)) + 2
        outs = [np.empty(side * side * 3 // 2 + 4096, np.uint8) for _ in group]     # a q100 file stays below 1.5 B per pixel
        try:
            ang, need, stat, sizes, fsz = ctx.correct_batch_files([loaded[t.task_id] for t in group], *params, outs=outs,
                                                                  jpeg_quality=jpeg_quality)
        except api.OicsError:
            leftovers += group
            continue
        for i, t in enumerate(group):
            if stat[i] != 0:
                leftovers.append(t)
                continue
            try:
                outs[i][:int(fsz[i])].tofile(t.output_path)
                result = "debatable" if need[i] else "finished"
            except OSError:
                result = "error"
            emit("task_completed", {"task_id": t.task_id, "output_path": t.output_path, "result": result})
    for t in leftovers:           # not decodable on the device: the caller's decoder, then the pixel path
        img = None
        if fallback_decode is not None and loaded[t.task_id] is not None:
            img = fallback_decode(loaded[t.task_id].tobytes())
        if img is None:
            emit("task_completed", {"task_id": t.task_id, "output_path": t.output_path, "result": "error"})
            continue
        done = []
        run_tasks(ctx, [Task(t.task_id, np.ascontiguousarray(img), t.output_path)], *params,
                  emit=lambda e, p: done.append((e, p)),
                  write=lambda path, image: np.frombuffer(ctx.jpeg_encode(image, jpeg_quality), np.uint8).tofile(path))
        for e, p in done:
            if e == "task_completed":
                emit(e, p)
