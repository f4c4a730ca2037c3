"""Inert stand-in for timm (only `timm.models.layers.Mlp` is imported, A2M-only)."""
