"""Structural types for editors (reference: python/vulkan_forge/types.py)."""
from __future__ import annotations

from typing import Any, Protocol


class SupportsPNG(Protocol):
    """Anything with ``render_png(path)`` -- Renderer, TerrainSpike, Scene."""

    def render_png(self, path: str) -> Any: ...
