"""Pure-Python utilities: data (IDX reader, synthetic MNIST-shaped generator), checkpoints in the
reference client's layout, configuration, logging/metrics, timing and clock sampling."""
