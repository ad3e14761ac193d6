"""Drop-in for the reference package `simple_knn` (KNN/): `from simple_knn._C import distCUDA2`
(TS/geometry/surfel_base.py:20, gaussian_base.py:25) keeps working unchanged."""
