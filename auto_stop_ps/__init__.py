"""auto_stop_ps variant of the trainer (ps tasks exit once every worker has reported in)."""
