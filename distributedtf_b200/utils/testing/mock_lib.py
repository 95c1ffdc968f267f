"""Mock benchmark logger (reference ``official/utils/testing/mock_lib.py:23-36``)."""


class MockBenchmarkLogger:
    """Records every ``log_metric`` call for inspection in tests."""

    def __init__(self):
        self.logged_metric = []

    def log_metric(self, name, value, unit=None, global_step=None, extras=None):
        self.logged_metric.append({"name": name, "value": float(value), "unit": unit,
                                   "global_step": global_step, "extras": extras})
