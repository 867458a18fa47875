"""simple_knn._C replacement (reference simple-knn/ext.cpp:4-6): distCUDA2 implemented in rain_b200."""
from rain_b200._C import distCUDA2  # noqa: F401
