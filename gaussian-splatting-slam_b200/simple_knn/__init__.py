"""Drop-in for the `simple_knn` package (submodules/simple-knn); use `from simple_knn._C import distCUDA2`
exactly as scene/gaussian_model.py:20 does."""
