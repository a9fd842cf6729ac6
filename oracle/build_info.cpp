// what the reference's cmake would generate from common/build-info.cpp.in (R/common/build-info.cpp.in:1-4)
int LLAMA_BUILD_NUMBER = 0;
char const *LLAMA_COMMIT = "018b7af";
char const *LLAMA_COMPILER = "gcc (oracle/Makefile)";
char const *LLAMA_BUILD_TARGET = "x86_64-linux-gnu";
