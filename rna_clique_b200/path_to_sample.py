"""Sample name of a top-genes FASTA path (reference path_to_sample.py:6-31)."""
import re
from pathlib import Path

sample_re = re.compile(r"(.*)_top\.fasta")


class PathToSampleError(Exception):
    def __init__(self, message, path):
        super().__init__(message)
        self.path = path


def path_to_sample(x) -> str:
    res = sample_re.search(Path(x).name)
    if not res:
        raise PathToSampleError(f"Unable to parse sample name from path {x} using regular expression {sample_re.pattern}.", x)
    return res.group(1)


def dict_path_to_sample(d):
    def inner(p):
        try:
            return d[Path(p)]
        except KeyError:
            raise PathToSampleError(f"Cannot get sample name for path {p}.", p) from None
    return inner
