"""Multi-GPU runtime: symmetric-heap engine, GPU worker, DDP communication hook."""
