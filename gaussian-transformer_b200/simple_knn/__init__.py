"""Drop-in for the reference's `simple_knn` package (submodules/simple-knn): `from simple_knn._C import distCUDA2`
(scene/gaussian_model.py:20,134).  Backed by libgsrast.so (gsr_knn_dist2); no CPU fallback."""
