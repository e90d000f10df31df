from .weight_utils import initialize_dummy_weights, load_dummy_model, seed_integer_params

__all__ = ["initialize_dummy_weights", "load_dummy_model", "seed_integer_params"]
