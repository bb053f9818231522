"""Model families used by the data-parallel benchmarks."""
