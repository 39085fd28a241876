"""Drop-in for the reference's `simple_knn` package (scene/gaussian_model.py:20: `from simple_knn._C import distCUDA2`)."""
