/* Stand-in for the header the reference's CMake generates from src/fimex_config.h.in.
 * Written for oracle/Makefile; not a copy of any generated file. */
#ifndef FIMEX_CONFIG_H
#define FIMEX_CONFIG_H 1
#define FIMEX_DATADIR "/nonexistent"
#define FIMEX_IO_PLUGINS_PATH "/nonexistent"
#define FIMEX_IO_PLUGINS_VERSION ""
#define HAVE_PROJ_H 1
#endif
