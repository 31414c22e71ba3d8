"""Test-only stand-in for matplotlib (absent from this image)."""
