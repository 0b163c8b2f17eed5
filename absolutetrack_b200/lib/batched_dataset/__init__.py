"""GPU side of the reference's lib/batched_dataset (config 3, torch_data sequences)."""
