from . import env_checker, seeding  # noqa: F401
