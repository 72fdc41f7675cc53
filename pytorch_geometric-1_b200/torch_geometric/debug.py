__debug_flag__ = {'enabled': False}


def is_debug_enabled():
    return __debug_flag__['enabled']


def set_debug_enabled(mode):
    __debug_flag__['enabled'] = mode


class debug(object):
    def __init__(self):
        self.prev = is_debug_enabled()

    def __enter__(self):
        set_debug_enabled(True)

    def __exit__(self, *args):
        set_debug_enabled(self.prev)
        return False


class set_debug(object):
    def __init__(self, mode):
        self.prev = is_debug_enabled()
        set_debug_enabled(mode)

    def __enter__(self):
        pass

    def __exit__(self, *args):
        set_debug_enabled(self.prev)
        return False
