"""gsb200: host-side mirror of the reference interfaces on the rasterizer / loss / densify hot path.

  gsb200.raster   tensor-level forward / backward / mark_visible (single camera and camera batches)
  gsb200.loss     fused L1 + SSIM (calc_ssim drop-in, external.py:68-76)
  gsb200.densify  densification statistics and clone / split / prune (external.py:113-314)
  gsb200.dist     camera sharding over ranks + one NCCL allreduce of gradients and statistics
  gsb200.scene    seeded synthetic Gaussians and cameras in the reference's conventions
All compute goes through libgsb200.so (include/gsb200.h); nothing here falls back to PyTorch or the CPU.
"""
