from .linear_stack import MODEL_SHAPES, AWQLinearStack, forward_tp_emulated, linear_shapes

__all__ = ["MODEL_SHAPES", "AWQLinearStack", "forward_tp_emulated", "linear_shapes"]
