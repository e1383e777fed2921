# Convenience targets; everything also works as plain python commands (see README.md).
PY ?= python

build:
	$(PY) __graft_entry__.py

test-cpu: build
	$(PY) -m pytest tests -x -q -m "not gpu"

test-gpu: build
	$(PY) -m pytest tests -x -q -m gpu

smoke: build
	$(PY) __graft_entry__.py smoke

bench: build
	$(PY) bench.py

golden:
	$(PY) tests/golden/make_golden.py

.PHONY: build test-cpu test-gpu smoke bench golden
