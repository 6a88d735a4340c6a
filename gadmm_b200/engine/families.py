"""Remaining algorithm families on the GPU (filled in by later modules)."""
