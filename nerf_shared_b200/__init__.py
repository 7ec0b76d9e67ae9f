"""nerf_shared on B200: the reference's Python API (nerf / render_utils / utils / config_parser)
over hand-written sm_100a CUDA kernels (libnerf_b200.so, C ABI in include/nerf_b200.h)."""
__all__ = ["config_parser", "nerf", "render_utils", "utils"]
__version__ = "0.1.0"
