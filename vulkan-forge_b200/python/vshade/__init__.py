"""Legacy alias package (reference: python/vshade/__init__.py:1-2)."""
from vulkan_forge import Renderer, __version__, render_triangle_png, render_triangle_rgba

__all__ = ["Renderer", "render_triangle_rgba", "render_triangle_png", "__version__"]
