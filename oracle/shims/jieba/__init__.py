def lcut(s):
    return list(s)
def initialize():
    pass
