"""No-op logging, call sites e.g. reference motion_primitive_generator.py:32,121."""
LOG_MODE_ERROR = -1
LOG_MODE_INFO = 1
LOG_MODE_DEBUG = 2
_active = set()


def write_message_to_log(message, mode=LOG_MODE_INFO):
    if mode in _active:
        print(message)


def write_log(*args):
    pass


def set_log_mode(mode):
    pass
