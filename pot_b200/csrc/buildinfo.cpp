// Build provenance of libb2ot.so: hash of the CUDA sources and headers the library was built from (Makefile: SRC_HASH).
#ifndef B2OT_SRC_HASH
#define B2OT_SRC_HASH "unknown"
#endif
extern "C" const char* b2ot_build_info(void) { return "src=" B2OT_SRC_HASH " built=" __DATE__ " " __TIME__; }
