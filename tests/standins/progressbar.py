"""Stand-in for the `progressbar2` package the reference's training scripts import (absent from this image)."""


def progressbar(iterable, *args, **kwargs):
    return iterable
