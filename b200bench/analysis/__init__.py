from .parse_metrics import collect_results, add_scaling_efficiency  # noqa: F401
