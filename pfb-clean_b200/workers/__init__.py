"""Device-side pieces of the reference workers (pfb/workers) that touch the cube between minor cycles."""
