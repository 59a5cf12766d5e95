"""Argument checks done in Python before the native module is reached.

Behavioural mirror of the reference's python/vulkan_forge/_validate.py:6-39: every failure is a
ValueError; sizes are capped at 8192, grids at 2..4096, PNG paths need a .png suffix and an
existing parent directory.
"""
from __future__ import annotations

import os
from pathlib import Path

MAX_TARGET_DIM = 8192
MIN_GRID, MAX_GRID = 2, 4096


def _to_int(label: str, value) -> int:
    try:
        return int(value)
    except Exception as exc:
        raise ValueError(f"{label} must be an integer, got {type(value).__name__}") from exc


def size_wh(width, height):
    dims = (_to_int("width", width), _to_int("height", height))
    if min(dims) <= 0:
        raise ValueError("width and height must be > 0")
    if max(dims) > MAX_TARGET_DIM:
        raise ValueError(f"width/height must be <= {MAX_TARGET_DIM}")
    return dims


def grid(n) -> int:
    value = _to_int("grid", n)
    if value < MIN_GRID:
        raise ValueError(f"grid must be >= {MIN_GRID}")
    if value > MAX_GRID:
        raise ValueError(f"grid must be <= {MAX_GRID}")
    return value


def png_path(p) -> str:
    text = os.fspath(p) if isinstance(p, Path) else str(p)
    if not text.lower().endswith(".png"):
        raise ValueError("path must end with .png")
    folder = Path(text).resolve().parent
    if not folder.exists():
        raise ValueError(f"directory does not exist: {folder}")
    return text
