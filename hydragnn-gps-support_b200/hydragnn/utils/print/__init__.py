from .print_utils import print_distributed, print_master, setup_log, log, iterate_tqdm
