extern "C" int dtf_io_version(void) { return 1; }
