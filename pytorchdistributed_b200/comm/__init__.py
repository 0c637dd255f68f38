from .process_group import ddp_setup, destroy_process_group, is_main, local_rank, pick_backend, world

__all__ = ["ddp_setup", "destroy_process_group", "is_main", "local_rank", "pick_backend", "world"]
