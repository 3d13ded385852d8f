# Convenience front end of hpnn_b200/build.py (the reference uses autotools: ./autogen.sh &&
# ./configure [--enable-cuda ...] && make, INSTALL:16-150).  Everything is built in-tree for
# sm_100a; no GPU is needed to build.
PYTHON ?= python
PREFIX ?= /usr/local

.PHONY: all lib tools check check-gpu bench reference clean install pc

all: lib

lib tools:
	$(PYTHON) hpnn_b200/build.py -q

check: lib
	$(PYTHON) -m pytest tests -x -q -m "not gpu"

check-gpu: lib
	$(PYTHON) -m pytest tests -x -q -m gpu

bench: lib
	$(PYTHON) bench.py

reference:
	$(PYTHON) baseline/build_ref.py

pc:
	@sed -e 's|@prefix@|$(PREFIX)|g' -e 's|@version@|0.1.0|g' csrc/capi/libhpnn.pc.in > build/libhpnn.pc

install: lib pc
	install -d $(DESTDIR)$(PREFIX)/lib $(DESTDIR)$(PREFIX)/include/libhpnn $(DESTDIR)$(PREFIX)/bin $(DESTDIR)$(PREFIX)/lib/pkgconfig
	install -m 755 hpnn_b200/lib/libhpnn.so hpnn_b200/lib/libhpnn_core.so $(DESTDIR)$(PREFIX)/lib/
	install -m 644 include/libhpnn.h include/hpnn_core.h $(DESTDIR)$(PREFIX)/include/
	install -m 644 include/libhpnn/common.h include/libhpnn/ann.h include/libhpnn/snn.h include/libhpnn/unroll.def $(DESTDIR)$(PREFIX)/include/libhpnn/
	install -m 755 build/bin/train_nn build/bin/run_nn build/bin/pmnist build/bin/pdif $(DESTDIR)$(PREFIX)/bin/
	install -m 644 build/libhpnn.pc $(DESTDIR)$(PREFIX)/lib/pkgconfig/

clean:
	rm -rf build hpnn_b200/lib
