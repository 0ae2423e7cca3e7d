"""TEST INFRASTRUCTURE: import-only stub (matplotlib is absent; the reference imports it at module
scope in iit/tasks/ioi/ioi_config.py:19 and iit/utils/plotter.py:1 but the hot path never plots)."""
