# Top-level build, same targets as the reference's (/root/reference/Makefile:1-15):
#   make            the sm_100a library (nvcc) and the synthetic-read generator
#   make install    populates PREFIX (default bin/) with the reference's installed layout
#                   (/root/reference/Makefile:4-11): select_STR_reads (driver script), VERSION,
#                   select_STR_reads.so and fastq.so (extension modules under the reference's module
#                   names), plus libbaitstr_select.so (C ABI + embedded sm_100a code) and the
#                   Python package with the -c / -d diagnostics; merge_STR_reads (the reference's
#                   command line over bstr_merge_reads) and extend_STR_reads (over bstr_kmers_*).
PREFIX ?= bin
PYTHON ?= python3

all:
	$(MAKE) -C baitstr_b200/csrc

install: all
	mkdir -p $(PREFIX)/baitstr_b200
	[ "$(abspath $(PREFIX))" = "$(abspath bin)" ] || cp bin/select_STR_reads bin/merge_STR_reads bin/extend_STR_reads bin/VERSION $(PREFIX)/
	cp baitstr_b200/libbaitstr_select.so $(PREFIX)/
	cp baitstr_b200/*.py $(PREFIX)/baitstr_b200/
	$(PYTHON) baitstr_b200/ext/build_ext.py $(PREFIX) $(PREFIX)

clean:
	$(MAKE) -C baitstr_b200/csrc clean
	-rm -rf bin/baitstr_b200 bin/*.so

.PHONY: all install clean
