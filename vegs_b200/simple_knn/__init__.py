"""Drop-in replacement for VEGS's `simple_knn` package (submodules/simple-knn)."""
