from . import ops, schedule_utils  # noqa: F401
