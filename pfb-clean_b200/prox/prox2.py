"""pfb/prox/prox2.py:3"""


def prox2(x, gamma):
    return x / (1 + gamma)
