"""packages/lib/src/types.rs: ImageFormat, RotateClipStrategy."""
import enum


class ImageFormat(enum.IntEnum):          # types.rs:1-5
    JPEG = 0
    PNG = 1
    WEBP = 2


class RotateClipStrategy(enum.IntEnum):   # types.rs:7-11
    DEFAULT = 0
    CONTAIN = 1
