// Syntax-check stand-in for the un-vendored third-party gflags header.
#pragma once
#include <string>
#define DECLARE_bool(name) extern bool FLAGS_##name
#define DECLARE_int32(name) extern int FLAGS_##name
#define DECLARE_int64(name) extern long long FLAGS_##name
#define DECLARE_uint64(name) extern unsigned long long FLAGS_##name
#define DECLARE_double(name) extern double FLAGS_##name
#define DECLARE_string(name) extern std::string FLAGS_##name
