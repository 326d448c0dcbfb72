"""Kept entry points of the reference's hot-path scripts."""
