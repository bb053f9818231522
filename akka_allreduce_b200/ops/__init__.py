"""Hand-written sm_100a kernels and their launchers."""
