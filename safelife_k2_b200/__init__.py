"""
safelife_k2_b200 — B200-native batched SafeLife hot path (advance_board, execute_action,
points/performance, side-effect rollouts) behind the reference's own interfaces.

    from safelife_k2_b200 import speedups            # speedups.advance_board(board, spawn_prob)
    from safelife_k2_b200.safelife_game import SafeLifeGame, LevelPool, CellTypes
    from safelife_k2_b200.safelife_env import SafeLifeEnv
    from safelife_k2_b200.side_effects import side_effect_score

All compute is CUDA (libsafelife_b200.so, sm_100a) through the C ABI in include/safelife_b200.h.
There is no CPU fallback; importing the compute modules without the built library raises.
"""
__all__ = ["speedups", "safelife_game", "safelife_env", "side_effects", "env_wrappers", "benchmarking",
           "distributed"]
