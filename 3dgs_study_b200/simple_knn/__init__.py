"""Drop-in for the reference's `simple_knn` package (import: /root/reference/scene/gaussian_model.py:20)."""
