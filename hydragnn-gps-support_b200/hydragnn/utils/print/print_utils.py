"""Verbosity-gated logging (reference: hydragnn/utils/print/print_utils.py:42-107)."""
import logging
import os

from ..distributed import get_comm_size_and_rank


def print_master(*args, **kwargs):
    if get_comm_size_and_rank()[1] == 0:
        print(*args, **kwargs)


def print_distributed(verbosity_level, *args, **kwargs):
    """0: nothing, 1-2: rank 0 only, 3-4: every rank (prefixed)."""
    size, rank = get_comm_size_and_rank()
    if verbosity_level in (1, 2) and rank == 0:
        print(*args, **kwargs)
    elif verbosity_level >= 3:
        print(f"{rank}:", *args, **kwargs)


def iterate_tqdm(iterator, verbosity_level, *args, **kwargs):
    if verbosity_level in (2, 4) and get_comm_size_and_rank()[1] == 0:
        from tqdm import tqdm
        return tqdm(iterator, *args, **kwargs)
    return iterator


def setup_log(prefix, path="./logs/"):
    """./logs/<prefix>/run.log + console, rank-prefixed."""
    _, rank = get_comm_size_and_rank()
    os.makedirs(os.path.join(path, prefix), exist_ok=True)
    logger = logging.getLogger("hydragnn")
    logger.setLevel(logging.DEBUG)
    for h in list(logger.handlers):
        logger.removeHandler(h)
    fmt = logging.Formatter(f"%(levelname)s (rank {rank}): %(message)s")
    fh = logging.FileHandler(os.path.join(path, prefix, "run.log"))
    fh.setFormatter(fmt)
    logger.addHandler(fh)
    return logger


def log(*args, sep=" "):
    logging.getLogger("hydragnn").info(sep.join(map(str, args)))
