"""Inert stand-in for the MATLAB engine package (channel.py:1 imports it at module import)."""
