"oracle-ref-build\n"
