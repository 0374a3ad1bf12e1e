# Thin front end to build.py (the reference's Makefile builds only `riemann`, with an undefined
# $(CFLAGS) and a target whose source is missing: Makefile:1-9).  Everything is built in-tree for
# sm_100a; `make riemann` etc. keep the reference's habit of naming the program you want.
PYTHON ?= python

all:
	$(PYTHON) build.py

riemann cintegrate train4: all
	@ls -l bin/$@

reference:
	$(PYTHON) baseline/install.py

test:
	$(PYTHON) -m pytest tests -q -m "not gpu"

test-gpu:
	$(PYTHON) -m pytest tests -q -m gpu

bench:
	$(PYTHON) bench.py --gpus 1

clean:
	rm -rf build bin b200quad/lib/*.so

.PHONY: all riemann cintegrate train4 reference test test-gpu bench clean
