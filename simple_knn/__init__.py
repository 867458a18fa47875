"""Drop-in package for `from simple_knn._C import distCUDA2` (scene/gaussian_model.py:9 of sharonal10/rain)."""
