"""TEST INFRASTRUCTURE ONLY -- stand-in for trimesh (io/mesh.py:22)."""
