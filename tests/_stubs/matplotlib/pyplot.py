"""Test-only stand-in for matplotlib.pyplot."""
