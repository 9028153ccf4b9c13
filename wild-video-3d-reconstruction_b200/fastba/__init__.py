from .ba import BA, BA_sharded, neighbors, reproject, solve_system  # noqa: F401
