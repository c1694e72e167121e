/*
 * integration/sfx_suffixer_stubs.c -- symbols of src/match/sfx-suffixer.c that are NOT on the
 * suffixerator path (the k-mer streaming entry points used by `gt encseq2spm`, readjoiner and
 * `gt dev` tools, SURVEY.md section 2.2 "firstcodes").  The B200 build of the Sfxiterator does not
 * provide them; they exist only so that unrelated objects of libgenometools still link.
 */
#include <stdio.h>
#include <stdlib.h>

#define B200_STUB(name)                                                                  \
  void name(void)                                                                        \
  {                                                                                      \
    fprintf(stderr, "%s: not available in the B200 build of sfx-suffixer\n", #name);     \
    exit(EXIT_FAILURE);                                                                  \
  }

B200_STUB(getencseqkmers_twobitencoding)
B200_STUB(getencseqkmersinsertkmerwithoutspecial)
B200_STUB(gt_getencseqkmers_twobitencoding_slice)
B200_STUB(gt_kmercode_at_firstpos)
