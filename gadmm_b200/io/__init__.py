from .matio import (load_input_data, save_input_data, load_optimal_sol, save_optimal_sol,
                    load_text_dataset, save_curves, save_state, load_state)

__all__ = ["load_input_data", "save_input_data", "load_optimal_sol", "save_optimal_sol",
           "load_text_dataset", "save_curves", "save_state", "load_state"]
