registry = {}


def register(id, entry_point):
    registry[id] = entry_point
