"""Data-parallel machinery: symmetric arena, fused gradient reducer, fused optimizer, launcher."""
