// util.h — stand-in for the reference's unshipped helper header
// (/root/reference/namegensf.cu:2).  The reference needs read_binary()
// (ref:370-372) and, transitively, <cstdio>/<cstdlib>/<cstring>.
#pragma once
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>

void *read_binary(const char *filename, size_t *size);
