/* Hand-written config.h used ONLY to compile the unmodified uvaia reference (nearest.c path)
 * as the parity oracle.  The reference's autotools build is not runnable in this image
 * (no autoconf/automake/libtool, no lzma.h/bzlib.h), so this is a zlib-only configuration. */
#ifndef ORACLE_CONFIG_H
#define ORACLE_CONFIG_H
#define HAVE_ZLIB 1
#define HAVE_SYS_TIME_H 1
#define HAVE_SYS_TIMES_H 1
#define HAVE_UNISTD_H 1
#define PACKAGE "uvaia"
#define PACKAGE_NAME "uvaia"
#define PACKAGE_VERSION "2.0.2"
#define PACKAGE_STRING "uvaia 2.0.2"
#define VERSION "2.0.2"
#define SIZEOF_INT 4
#define SIZEOF_DOUBLE 8
#define SIZEOF_LONG_LONG_INT 8
#ifndef _GNU_SOURCE
#define _GNU_SOURCE 1
#endif
#endif
