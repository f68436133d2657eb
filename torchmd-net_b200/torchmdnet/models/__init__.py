"""Model names accepted by ``create_model`` (reference: torchmdnet/models/__init__.py)."""
__all__ = ["graph-network", "transformer", "equivariant-transformer", "tensornet"]
