"""Drop-in for the reference's pybind module simple_knn._C (thirdparty/simple-knn/ext.cpp:15-17)."""
from wildgs_slam_blackwell_b200.knn import distCUDA2  # noqa: F401
