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


class _HostBuffer:
    """Page-locked when the CUDA runtime can provide it, pageable otherwise (a buffer, not a compute path)."""

    def __init__(self, nbytes: int):
        self._pin = None
        try:
            self._pin = api.PinnedArray(nbytes)
            self.array = self._pin.array
        except api.OicsError:
            self.array = np.empty(nbytes, np.uint8)

    def close(self):
        if self._pin is not None:
            self._pin.close()
        self.array = None


def run_tasks(ctx: "api.Context", tasks: List[Task], projection_max_angle: int, projection_angle_step: float,
              projection_max_width: int, projection_max_height: int, hough_min_line_length: float,
              hough_max_line_gap: float, emit: Callable[[str, dict], None],
              write: Optional[Callable[[str, np.ndarray], None]] = None, chunk: int = 32) -> None:
    """Runs the tasks and emits the reference's events through `emit(event_name, payload)`.

    The tasks of one card size go through `oics_correct_batch` in chunks of `chunk` cards: `start_running_task`
    is emitted when a chunk is handed to the device and every `task_completed` of the chunk as soon as the chunk
    returns -- the reference emits per task as its thread finishes (task.rs:32-67); a chunk is the unit the
    pipeline completes.  Cards are gathered into, and deskewed cards returned through, page-locked buffers
    (oics_alloc_pinned) so the copies overlap the kernels."""
    by_shape = {}
    for t in tasks:
        by_shape.setdefault(tuple(t.card.shape), []).append(t)
    for shape, group in by_shape.items():
        if len(shape) != 3 or shape[2] != 3:
            for t in group:
                emit("start_running_task", {"task_id": t.task_id})
                emit("task_completed", {"task_id": t.task_id, "output_path": t.output_path, "result": "error"})
            continue
        h, w = shape[:2]
        side = int(np.ceil((w + h) * 0.70710678118654757)) + 2          # CONTAIN bound of any rotation
        n_chunk = max(1, min(chunk, len(group)))
        card_bytes, out_bytes = h * w * 3, side * side * 3
        pin_in = _HostBuffer(n_chunk * card_bytes)
        pin_out = _HostBuffer(n_chunk * out_bytes) if write else None
        try:
            for c0 in range(0, len(group), n_chunk):
                part = group[c0:c0 + n_chunk]
                cards = pin_in.array[:len(part) * card_bytes].reshape(len(part), h, w, 3)
                for i, t in enumerate(part):
                    emit("start_running_task", {"task_id": t.task_id})
                    cards[i] = t.card
                outs = [pin_out.array[i * out_bytes:(i + 1) * out_bytes] for i in range(len(part))] if write else None
                try:
                    ang, need, stat, sizes = ctx.correct_batch(cards, projection_max_angle, projection_angle_step,
                                                               projection_max_width, projection_max_height,
                                                               hough_min_line_length, hough_max_line_gap, outs=outs)
                except api.OicsError:
                    for t in part:
                        emit("task_completed", {"task_id": t.task_id, "output_path": t.output_path, "result": "error"})
                    continue
                for i, t in enumerate(part):
                    if stat[i] != 0:
                        result = "error"                                            # task.rs:58-62: any Err
                    else:
                        result = "debatable" if need[i] else "finished"             # task.rs:47-57
                        if write:
                            dw, dh = int(sizes[i][0]), int(sizes[i][1])
                            write(t.output_path, outs[i][:dw * dh * 3].reshape(dh, dw, 3))
                    emit("task_completed", {"task_id": t.task_id, "output_path": t.output_path, "result": result})
        finally:
            pin_in.close()
            if pin_out is not None:
                pin_out.close()


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
