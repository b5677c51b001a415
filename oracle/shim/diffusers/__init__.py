"""Minimal stand-in for the two `diffusers` symbols the reference DiT imports (test infrastructure only)."""
