/*
 * TEST INFRASTRUCTURE ONLY (oracle/): a 40-line driver that exposes two entry points of the
 * UNMODIFIED reference (genometools 1.6.7 under /root/reference) as a stand-alone binary:
 *
 *   gt_ref suffixerator <args>   -> gt_parseargsandcallsuffixerator (src/match/sfx-run.c:719),
 *                                   what `gt suffixerator` runs (src/tools/gt_suffixerator.c:22-26)
 *   gt_ref sfxmap <args>         -> the checker tool `gt dev sfxmap` (src/tools/gt_sfxmap.c)
 *   gt_ref packedindex <args>    -> `gt packedindex` (src/tools/gt_packedindex.c): `mkindex` builds the packed (FM) index
 *                                   THROUGH the Sfxiterator (src/match/eis-suffixerator-interface.c:254,400), the second
 *                                   in-tree caller of the interface the drop-in replaces
 *
 * It is compiled by oracle/Makefile against the reference sources where they lie; nothing of the
 * reference is copied into this repository. Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may execute the resulting oracle/_ref/gt_ref.
 */
#include <stdio.h>
#include <string.h>
#include "core/init_api.h"
#include "core/error_api.h"
#include "core/tool_api.h"
#include "match/sfx-run.h"
#include "tools/gt_sfxmap.h"
#include "tools/gt_packedindex.h"

int main(int argc, const char **argv)
{
  GtError *err;
  int had_err = 0;

  if (argc < 2) {
    fprintf(stderr, "usage: %s suffixerator|sfxmap|packedindex [options]\n", argv[0]);
    return 2;
  }
  gt_lib_init();
  err = gt_error_new();
  gt_error_set_progname(err, argv[1]);
  if (strcmp(argv[1], "suffixerator") == 0) {
    had_err = gt_parseargsandcallsuffixerator(true, argc - 1, argv + 1, err);
  } else if (strcmp(argv[1], "sfxmap") == 0) {
    GtTool *tool = gt_sfxmap();
    had_err = gt_tool_run(tool, argc - 1, argv + 1, err);
    gt_tool_delete(tool);
  } else if (strcmp(argv[1], "packedindex") == 0) {
    GtTool *tool = gt_packedindex();
    had_err = gt_tool_run(tool, argc - 1, argv + 1, err);
    gt_tool_delete(tool);
  } else {
    fprintf(stderr, "%s: unknown subcommand %s\n", argv[0], argv[1]);
    had_err = 2;
  }
  if (gt_error_is_set(err)) {
    fprintf(stderr, "%s: error: %s\n", argv[1], gt_error_get(err));
    had_err = 1;
  }
  gt_error_delete(err);
  if (gt_lib_clean())
    return 2;
  return had_err ? 1 : 0;
}
