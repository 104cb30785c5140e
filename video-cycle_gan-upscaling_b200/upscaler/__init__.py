"""B200-native drop-in for the reference's `upscaler` package (upscaling/upscaler/{model,data}.py).

`upscaler.model` keeps the reference's builder / loss / compile-helper names; everything below it runs as hand-written
sm_100a CUDA kernels reached through the C ABI in include/b2srgan.h (libb2srgan.so).
"""
__all__ = ["model", "data"]
