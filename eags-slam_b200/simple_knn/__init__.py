"""Drop-in for the un-vendored ``simple_knn`` pip package (reference: ``requirements.txt`` /
``src/entities/gaussian_model.py:17``): ``from simple_knn._C import distCUDA2`` keeps working."""
