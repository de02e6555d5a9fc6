"""see oracle/shims/matplotlib/__init__.py: imported by data_fm.py:1, never called."""
