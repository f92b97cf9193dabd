/*
 * splinter_gpu_search -- the GPU `search` module (splinter_cli_cmd_search_gpu.c) as a stand-alone command, for
 * hosts that want the search without the rest of the reference's CLI:
 *
 *   splinter_gpu_search --use STORE <query>|- [--file PATH] [--json] [--limit N] [--distance F]
 *                       [--similarity F] [--bloom HEXMASK] [--regex PATTERN]
 *
 * It supplies the two things the module expects from the reference's CLI framework -- the session record
 * `thisuser` (splinter_cli_main.c:329-346) and cli_show_key_type (splinter_cli_util.c:128-139) -- connects to
 * the store like the dispatcher's --use does (splinter_cli_main.c:752-756) and hands the remaining arguments
 * to cmd_search.  Links the reference's own libsplinter.so, which stays the owner of the store format.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "splinter_cli.h"

cli_user_t thisuser;

char *cli_show_key_type(unsigned short flags) {
    static const struct { unsigned short bit; const char *name; } names[] = {
        { SPL_SLOT_TYPE_BIGINT, "SPL_SLOT_TYPE_BIGINT" },   { SPL_SLOT_TYPE_BIGUINT, "SPL_SLOT_TYPE_BIGUINT" },
        { SPL_SLOT_TYPE_BINARY, "SPL_SLOT_TYPE_BINARY" },   { SPL_SLOT_TYPE_IMGDATA, "SPL_SLOT_TYPE_IMGDATA" },
        { SPL_SLOT_TYPE_VARTEXT, "SPL_SLOT_TYPE_VARTEXT" }, { SPL_SLOT_TYPE_AUDIO, "SPL_SLOT_TYPE_AUDIO" },
        { SPL_SLOT_TYPE_JSON, "SPL_SLOT_TYPE_JSON" },       { SPL_SLOT_TYPE_VOID, "SPL_SLOT_TYPE_VOID" } };
    for (size_t i = 0; i < sizeof names / sizeof names[0]; i++)
        if (flags & names[i].bit) return (char *)names[i].name;
    return (char *)"UNNAMED";
}

int main(int argc, char **argv) {
    const char *store = getenv("SPLINTER_DEFAULT_STORE");
    char **sargv = (char **)calloc((size_t)argc + 1, sizeof *sargv);
    int sargc = 0;
    sargv[sargc++] = (char *)"search";
    for (int i = 1; i < argc; i++) {
        if ((!strcmp(argv[i], "--use") || !strcmp(argv[i], "-u")) && i + 1 < argc) store = argv[++i];
        else sargv[sargc++] = argv[i];
    }
    if (store && splinter_open(store) == 0) {
        thisuser.store_conn = 1;
        strncpy(thisuser.store, store, sizeof thisuser.store - 1);
    }
    const int rc = cmd_search(sargc, sargv);
    fflush(stdout);
    free(sargv);
    return rc == 0 ? 0 : 1;
}
