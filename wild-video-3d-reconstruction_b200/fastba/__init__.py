from .ba import BA, neighbors, reproject, solve_system  # noqa: F401
