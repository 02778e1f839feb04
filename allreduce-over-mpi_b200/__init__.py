"""flextree-b200 package root (imported as ``flextree``; see ../flextree/__init__.py)."""
