"""Host mirror of the Tauri batch path (packages/app/src-tauri/src/task.rs:19-68 and
thread_pool.rs): the reference fans `add_task` calls out to one OS thread per card, each
emitting `start_running_task`, running omr::correct_default and emitting `task_completed`
with result "finished" | "debatable" | "error".  Here the queued tasks of one parameter set
run through `oics_correct_batch` (upload / estimate + deskew / download pipeline on the GPU)
and the same events are emitted, in task order, with the same payload fields.

Decoding the input file and encoding the output (imread / imwrite, omr.rs:349,447-448) stay
with the caller: a task carries the decoded BGR card, `write(output_path, image)` is the
caller's encoder (None = do not write).
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
