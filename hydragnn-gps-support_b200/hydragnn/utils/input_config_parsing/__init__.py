from .config_utils import (
    update_config,
    update_config_minmax,
    get_log_name_config,
    save_config,
    merge_config,
)
