"""Data-parallel integration: communicator bootstrap over torch.distributed and the DDP comm hook."""
