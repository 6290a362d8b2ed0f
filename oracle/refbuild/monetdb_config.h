/* Hand-written build configuration for compiling the reference GDK (libbat)
 * directly with gcc on x86-64 Linux/glibc, without running the reference's
 * cmake build.  This file is ours (test infrastructure): it only states facts
 * about this platform that the reference sources ask for via HAVE_* macros
 * (see /root/reference/monetdb_config.h.in for the list of knobs). */
#ifndef MT_SEEN_MONETDB_CONFIG_H
#define MT_SEEN_MONETDB_CONFIG_H 1

#ifndef _GNU_SOURCE
#define _GNU_SOURCE 1
#endif
#include <stdlib.h>
#include <sys/types.h>
#include <assert.h>
#include <errno.h>
#include <inttypes.h>
#include <stdarg.h>
#include <stdbool.h>
#include <stddef.h>
#include <stdio.h>
#include <string.h>
#include <strings.h>

#define HAVE_SYS_TYPES_H 1
#define HAVE_DLFCN_H 1
#define HAVE_FCNTL_H 1
#define HAVE_NETDB_H 1
#define HAVE_POLL_H 1
#define HAVE_PWD_H 1
#define HAVE_STRINGS_H 1
#define HAVE_SYS_FILE_H 1
#define HAVE_SYS_IOCTL_H 1
#define HAVE_SYS_MMAN_H 1
#define HAVE_SYS_PARAM_H 1
#define HAVE_SYS_RANDOM_H 1
#define HAVE_SYS_RESOURCE_H 1
#define HAVE_SYS_TIME_H 1
#define HAVE_SYS_TIMES_H 1
#define HAVE_SYS_UIO_H 1
#define HAVE_SYS_UN_H 1
#define HAVE_TERMIOS_H 1
#define HAVE_UNISTD_H 1
#define HAVE_SEMAPHORE_H 1
#define HAVE_GETOPT_H 1
#define HAVE_STDATOMIC_H 1
#define HAVE_DIRENT_H 1
#define HAVE_SYS_SOCKET_H 1
#define HAVE_GETTIMEOFDAY 1
#define HAVE_SYS_STAT_H 1
#define HAVE_FDATASYNC 1
#define HAVE_ACCEPT4 1
#define HAVE_ASCTIME_R 1
#define HAVE_CLOCK_GETTIME 1
#define HAVE_CTIME_R 1
#define HAVE_FALLOCATE 1
#define HAVE_FCNTL 1
#define HAVE_FSYNC 1
#define HAVE_FTIME 1
#define HAVE_GETENTROPY 1
#define HAVE_GETLOGIN 1
#define HAVE_GETOPT_LONG 1
#define HAVE_GETRLIMIT 1
#define HAVE_GETUID 1
#define HAVE_GMTIME_R 1
#define HAVE_LOCALTIME_R 1
#define HAVE_STRERROR_R 1
#define HAVE_LOCKF 1
#define HAVE_MADVISE 1
#define HAVE_MREMAP 1
#define HAVE_NANOSLEEP 1
#define HAVE_NL_LANGINFO 1
#define HAVE_PIPE2 1
#define HAVE_POLL 1
#define HAVE_POPEN 1
#define HAVE_POSIX_FADVISE 1
#define HAVE_POSIX_FALLOCATE 1
#define HAVE_POSIX_MADVISE 1
#define HAVE_SETSID 1
#define HAVE_SOCKATMARK 1
#define HAVE_SHUTDOWN 1
#define HAVE_SIGACTION 1
#define HAVE_SIGLONGJMP 1
#define HAVE_STRPTIME 1
#define HAVE_SYSCONF 1
#define HAVE_TIMES 1
#define HAVE_TM_GMTOFF 1
#define HAVE_UNAME 1
#define HAVE_PTHREAD_KILL 1
#define HAVE_PTHREAD_MUTEX_TIMEDLOCK 1
#define HAVE_PTHREAD_SETNAME_NP 1
#define HAVE_PTHREAD_SIGMASK 1
#define HAVE_GETOPT 1
#define HAVE_PTHREAD_H 1
#define HAVE_GETADDRINFO 1
#define HAVE_SOCKLEN_T 1

#define MAPI_PORT 50000
#define MAPI_PORT_STR "50000"
#define DIR_SEP '/'
#define PATH_SEP ':'
#define DIR_SEP_STR "/"
#define SO_PREFIX "lib"
#define SO_EXT ".so"
#define BINDIR "/usr/local/bin"
#define LIBDIR "/usr/local/lib"
#define LOCALSTATEDIR "/usr/local/var"

#define MONETDB_VERSION "11.52.0"
#define MONETDB_VERSION_MAJOR 11
#define MONETDB_VERSION_MINOR 52
#define MONETDB_VERSION_PATCH 0
#define GDK_VERSION "29.0.0"
#define GDK_VERSION_MAJOR 29
#define GDK_VERSION_MINOR 0
#define GDK_VERSION_PATCH 0
#define MAPI_VERSION "27.0.0"
#define MAPI_VERSION_MAJOR 27
#define MAPI_VERSION_MINOR 0
#define MAPI_VERSION_PATCH 0
#define MONETDB5_VERSION "37.0.0"
#define MONETDB5_VERSION_MAJOR 37
#define MONETDB5_VERSION_MINOR 0
#define MONETDB5_VERSION_PATCH 0
#define MONETDBE_VERSION "27.0.0"
#define MONETDBE_VERSION_MAJOR 27
#define MONETDBE_VERSION_MINOR 0
#define MONETDBE_VERSION_PATCH 0
#define STREAM_VERSION "27.0.0"
#define STREAM_VERSION_MAJOR 27
#define STREAM_VERSION_MINOR 0
#define STREAM_VERSION_PATCH 0
#define SQL_VERSION "15.0.0"
#define SQL_VERSION_MAJOR 15
#define SQL_VERSION_MINOR 0
#define SQL_VERSION_PATCH 0
#define HOST "x86_64-pc-linux-gnu"
#define MONETDB5_PASSWDHASH "SHA512"
#define MONETDB5_PASSWDHASH_TOKEN SHA512

#define SIZEOF_SIZE_T 8
#define SIZEOF_VOID_P 8
#define SIZEOF_CHAR 1
#define SIZEOF_SHORT 2
#define SIZEOF_INT 4
#define SIZEOF_LONG 8
#define SIZEOF_LONG_LONG 8
#define SIZEOF_DOUBLE 8
#define SIZEOF_WCHAR_T 4
#define HAVE_LONG_LONG 1
#define SIZEOF___INT128 16
#define SIZEOF___INT128_T 16
#define SIZEOF___UINT128_T 16
#define HAVE___INT128 1
#define HAVE___INT128_T 1
#define HAVE___UINT128_T 1
#define HAVE_HGE 1
typedef __int128 hge;
typedef unsigned __int128 uhge;
#define SIZEOF_HGE SIZEOF___INT128

#if defined(__USE_GNU) && defined(HAVE_STRERROR_R)
#define STRERROR_R_CHARP 1
#endif
#define SOCKET int
#define closesocket close
#define ioctlsocket ioctl
#ifndef _In_z_
#define _In_z_
#endif
#ifndef _Printf_format_string_
#define _Printf_format_string_
#endif
#define LIB_STARTUP_FUNC(f) \
	static void f(void) __attribute__((__constructor__)); \
	static void f(void)
#endif
