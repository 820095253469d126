// elaina-exec <conf.json> -- same command line as the reference executable (main.cpp:10-18), backed by libwost.so.
#include <cstdio>

#include "../../include/wost.h"

int main(int argc, char **argv)
{
    if (argc < 2)
    {
        fprintf(stderr, "usage: %s <conf.json>\n", argv[0]);
        return 1;
    }
    const wost_status rc = wost_run_config(argv[1]);
    if (rc != WOST_OK)
    {
        fprintf(stderr, "elaina-exec: error %d: %s\n", (int)rc, wost_last_error());
        return 1;
    }
    return 0;
}
