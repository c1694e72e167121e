/*
 * integration/sfx_suffixer_b200.c -- the B200 build of GenomeTools' Sfxiterator.
 *
 * A from-scratch implementation of the interface declared in src/match/sfx-suffixer.h:33-71
 * (gt_Sfxiterator_new_withadditionalvalues / _new / _next / _longest / _bcktab2file / _delete)
 * that is compiled INSTEAD of src/match/sfx-suffixer.c when libgenometools is built for B200.
 * Everything around it is the unmodified reference: option parser, GtEncseq, GtSuffixsortspace,
 * the .suf/.lcp/.llv/.bck/.prj writers.  The CPU loops of the original (k-mer counting, bucket
 * insertion, gt_sortallbuckets and every sorter below it, the LCP bookkeeping) are replaced by
 * calls into the C ABI of include/gtsfx_b200.h; there is no CPU fallback.
 *
 * integration/Makefile links it, together with the unmodified reference objects, into
 * oracle/_ref/gt_b200 (test infrastructure: the drop-in check of tests/test_gpu_dropin.py).
 * Only reference HEADERS are included; no reference source is copied.
 */
#include <limits.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "core/alphabet_api.h"
#include "core/encseq.h"
#include "core/error_api.h"
#include "core/fa_api.h"
#include "core/intbits.h"
#include "core/logger_api.h"
#include "core/ma_api.h"
#include "core/timer_api.h"
#include "core/unused_api.h"
#include "core/xansi_api.h"
#include "match/sfx-lcpvalues.h"
#include "match/sfx-strategy.h"
#include "match/sfx-suffixer.h"
#include "match/sfx-suffixgetset.h"

#include "gtsfx_b200.h"

/* added to the B200 build of src/match/sfx-lcpvalues.c (integration/sfx_lcpvalues_b200.c) */
void gt_Outlcpinfo_b200_writepart(GtOutlcpinfo *outlcpinfo, const uint8_t *lcp, GtUword width,
                                  const void *llv, GtUword nllv, double lcptabsum, GtUword maxbranchdepth);
bool gt_Outlcpinfo_b200_wantsdist(const GtOutlcpinfo *outlcpinfo);
void gt_Outlcpinfo_b200_distadd(GtOutlcpinfo *outlcpinfo, const uint8_t *lcp, GtUword from, GtUword to,
                                const void *llv, GtUword nllv, GtUword partoffset, GtUword *llvidx);

#define B200_MAXDEV 16

struct Sfxiterator
{
  gtb200_multi *dev;               /* one or several GPUs (GTB200_DEVICES) behind one handle */
  void *pinned_suf, *pinned_lcp;   /* host buffers page-locked for overlapped copies */
  const GtEncseq *encseq;          /* borrowed */
  GtOutlcpinfo *outlcpinfo;        /* borrowed, may be NULL */
  FILE *outfpbcktab;               /* written in _delete when parts > 1, like the reference */
  GtSuffixsortspace *suffixsortspace;
  GtTimer *sfxprogress;
  GtLogger *logger;
  bool suftabuint, bck_written, pin;
  unsigned int prefixlength;
  uint32_t *leftborder, *countspecialcodes, *distpfxidx;
  gtb200_countinfo ci;
  uint64_t *mincodes, *maxcodes, *widths;
  int numofparts, part;
  GtUword totallength, bufentries, tail_done, longest;
  bool longest_defined;
  uint8_t *lcpbuf;
  gtb200_largelcp *llvbuf;
  GtUword llvcap;
};

/* GTB200_DEVICES=0,1,4-7 lists the GPUs every part is spread over (north_star: "each of the 1/2/4/8 GPUs ...
   sorts and LCP-annotates its own disjoint bucket ranges"); GTB200_DEVICE=k is the single-GPU form. */
static int b200_devices(int *devs)
{
  const char *env = getenv("GTB200_DEVICES");
  int n = 0;

  if (env == NULL || *env == '\0')
  {
    env = getenv("GTB200_DEVICE");
    devs[0] = env != NULL ? atoi(env) : 0;
    return 1;
  }
  n = gtb200_parse_devices(env, devs, B200_MAXDEV);
  if (n == 0)
  {
    devs[0] = 0;
    n = 1;
  }
  return n;
}

static void *b200_rawptr(Sfxiterator *sfi)
{
  /* partoffset and bucketleftidx stay 0, so index 0 is the start of the buffer that
     gt_suffixsortspace_to_file (src/match/sfx-suffixgetset.c:462-478) writes from */
  if (sfi->suftabuint)
  {
    return (void *) gt_suffixsortspace_getptr_uint32(sfi->suffixsortspace, 0);
  }
  return (void *) gt_suffixsortspace_getptr_ulong(sfi->suffixsortspace, 0);
}

static void b200_write_padded(FILE *fp, const void *ptr, size_t bytes)
{
  /* gt_mapspec_write: every table is followed by zero bytes up to a multiple of 8
     (src/core/mapspec.c:350-365) */
  static const char zeros[8] = {0};
  if (bytes > 0)
  {
    gt_xfwrite(ptr, (size_t) 1, bytes, fp);
  }
  if (bytes % 8 != 0)
  {
    gt_xfwrite(zeros, (size_t) 1, 8 - bytes % 8, fp);
  }
}

static void b200_bcktab_to_file(Sfxiterator *sfi, FILE *fp)
{
  /* table order of assignbcktabmapspecification (src/match/bcktab.c:519-570) */
  b200_write_padded(fp, sfi->leftborder, sizeof (uint32_t) * (size_t) (sfi->ci.numofallcodes + 1));
  b200_write_padded(fp, sfi->countspecialcodes, sizeof (uint32_t) * (size_t) sfi->ci.numofspecialcodes);
  if (sfi->prefixlength > 2U)
  {
    b200_write_padded(fp, sfi->distpfxidx, sizeof (uint32_t) * (size_t) sfi->ci.numofdistpfxidx);
  }
  sfi->bck_written = true;
}

int gt_Sfxiterator_delete(Sfxiterator *sfi, GT_UNUSED GtError *err)
{
  if (sfi == NULL)
  {
    return 0;
  }
  if (sfi->outfpbcktab != NULL && !sfi->bck_written && sfi->leftborder != NULL)
  {
    /* parts > 1: the reference writes .bck here (src/match/sfx-suffixer.c:560-577) */
    b200_bcktab_to_file(sfi, sfi->outfpbcktab);
    gt_fa_fclose(sfi->outfpbcktab);
  }
  if (sfi->pinned_suf != NULL)
  {
    gtb200_host_unregister(sfi->pinned_suf);
  }
  if (sfi->pinned_lcp != NULL)
  {
    gtb200_host_unregister(sfi->pinned_lcp);
  }
  if (sfi->dev != NULL)
  {
    gtb200_multi_close(sfi->dev);
  }
  if (sfi->suffixsortspace != NULL)
  {
    gt_suffixsortspace_delete(sfi->suffixsortspace, false);
  }
  gt_free(sfi->leftborder);
  gt_free(sfi->countspecialcodes);
  gt_free(sfi->distpfxidx);
  gt_free(sfi->mincodes);
  gt_free(sfi->maxcodes);
  gt_free(sfi->widths);
  gt_free(sfi->lcpbuf);
  gt_free(sfi->llvbuf);
  gt_free(sfi);
  return 0;
}

Sfxiterator *gt_Sfxiterator_new_withadditionalvalues(const GtEncseq *encseq,
                                                     GtReadmode readmode,
                                                     unsigned int prefixlength,
                                                     unsigned int numofparts,
                                                     GtUword maximumspace,
                                                     void *voidoutlcpinfo,
                                                     FILE *outfpbcktab,
                                                     const Sfxstrategy *sfxstrategy,
                                                     GtTimer *sfxprogress,
                                                     GT_UNUSED bool withprogressbar,
                                                     GtLogger *logger,
                                                     GtError *err)
{
  Sfxiterator *sfi;
  gtb200_seqdesc seq;
  gtb200_range *ranges = NULL;
  GtUchar *encoded = NULL;
  GtUword largest = 0;
  unsigned int i;
  int p, devs[B200_MAXDEV], ndev;

  gt_error_check(err);
  if (sfxstrategy != NULL && (sfxstrategy->spmopt_minlength > 0 || sfxstrategy->compressedoutput))
  {
    gt_error_set(err, "the B200 build of the suffixerator supports neither -spmopt nor -compressedoutput");
    return NULL;
  }
  sfi = gt_calloc((size_t) 1, sizeof *sfi);
  sfi->encseq = encseq;
  sfi->outlcpinfo = (GtOutlcpinfo *) voidoutlcpinfo;
  sfi->outfpbcktab = outfpbcktab;
  sfi->sfxprogress = sfxprogress;
  sfi->logger = logger;
  sfi->prefixlength = prefixlength;
  sfi->suftabuint = sfxstrategy != NULL && sfxstrategy->suftabuint;
  sfi->totallength = gt_encseq_total_length(encseq);
  if (sfi->suftabuint && sfi->totallength > (GtUword) UINT_MAX)
  {
    sfi->suftabuint = false;
  }
  memset(&seq, 0, sizeof seq);
  seq.numofchars = gt_alphabet_num_of_chars(gt_encseq_alphabet(encseq));
  seq.totallength = sfi->totallength;
  seq.memory = GTB200_MEM_HOST;
  seq.readmode = (int) readmode;   /* GtReadmode values are the C ABI's (core/readmode_api.h:24-27) */
  /* a mirrored encseq (-mirrored) stores only the forward half: its logical sequence is read through the
     extractor below (src/core/encseq.c, gt_encseq_extract_encoded handles the virtual reverse complement) */
  if (seq.numofchars == 4U && gt_encseq_has_twobitencoding(encseq) && !gt_encseq_is_mirrored(encseq))
  {
    GtUword k = 0, nr = gt_encseq_realspecialranges(encseq);
    seq.twobitencoding = (const uint64_t *) gt_encseq_twobitencoding_export(encseq);
    seq.numofwords = gt_unitsoftwobitencoding(sfi->totallength);
    ranges = gt_malloc(sizeof *ranges * (size_t) (nr + 1));
    if (gt_encseq_has_specialranges(encseq))
    {
      GtRange range;
      GtSpecialrangeiterator *sri = gt_specialrangeiterator_new(encseq, true);
      while (gt_specialrangeiterator_next(sri, &range))
      {
        if (k > 0 && ranges[k-1].end == range.start)   /* iterator may split a run at page borders */
        {
          ranges[k-1].end = range.end;
        } else
        {
          gt_assert(k <= nr);
          ranges[k].start = range.start;
          ranges[k++].end = range.end;
        }
      }
      gt_specialrangeiterator_delete(sri);
    }
    seq.specialranges = ranges;
    seq.numofspecialranges = k;
  } else
  {
    encoded = gt_malloc(sizeof *encoded * (size_t) (sfi->totallength + 1));
    if (sfi->totallength > 0)
    {
      gt_encseq_extract_encoded(encseq, encoded, 0, sfi->totallength - 1);
    }
    seq.encoded = encoded;
  }
  if (sfxprogress != NULL)
  {
    gt_timer_show_progress(sfxprogress, "counting prefix distribution", stdout);
  }
  ndev = b200_devices(devs);
  if (gtb200_multi_open(&sfi->dev, devs, ndev, &seq, prefixlength) != 0)
  {
    gt_error_set(err, "%s", gtb200_last_error());
    gt_free(ranges);
    gt_free(encoded);
    (void) gt_Sfxiterator_delete(sfi, NULL);
    return NULL;
  }
  gt_free(ranges);
  gt_free(encoded);
  /* numofallcodes = sigma^pl, numofspecialcodes = sigma^(pl-1), distpfxidx = sum_{i=1}^{pl-2} sigma^i */
  {
    uint64_t allcodes = 1, speccodes = 1, d = 0, q = 1;
    for (i = 0; i < prefixlength; i++) allcodes *= seq.numofchars;
    for (i = 0; i + 1 < prefixlength; i++) speccodes *= seq.numofchars;
    for (i = 1; i + 1 < prefixlength; i++) { q *= seq.numofchars; d += q; }
    sfi->leftborder = gt_malloc(sizeof (uint32_t) * (size_t) (allcodes + 1));
    sfi->countspecialcodes = gt_malloc(sizeof (uint32_t) * (size_t) speccodes);
    sfi->distpfxidx = gt_malloc(sizeof (uint32_t) * (size_t) (d + 1));
  }
  if (gtb200_multi_count(sfi->dev, sfi->leftborder, sfi->countspecialcodes, sfi->distpfxidx, &sfi->ci) != 0)
  {
    gt_error_set(err, "%s", gtb200_last_error());
    (void) gt_Sfxiterator_delete(sfi, NULL);
    return NULL;
  }
  gt_logger_log(logger, "largest bucket size=" GT_WU, (GtUword) sfi->ci.largestbucket);
  if (numofparts == 0)
  {
    numofparts = 1U;
  }
  {
    /* The number of parts follows from memory like gt_suftabparts_fit_memlimit does it for the reference
       (src/match/sfx-partssuf.c:449-527): -memlimit bounds the host side (8+1 bytes per suffix of a part next to
       the bucket table), the free HBM of the smallest GPU bounds the device side (sequence + bucket tables resident,
       about 40 bytes per suffix of a device's share of a part).  More parts never change the output. */
    const double nsort = (double) sfi->ci.numofsuffixestosort;
    const double tablebytes = 4.0 * ((double) sfi->ci.numofallcodes + (double) sfi->ci.numofspecialcodes
                                     + (double) sfi->ci.numofdistpfxidx);
    uint64_t freeb = 0, totalb = 0;
    unsigned int need = 1U;

    if (maximumspace > 0)
    {
      const double room = (double) maximumspace - tablebytes;
      const double per = (sfi->suftabuint ? 4.0 : 8.0) + (sfi->outlcpinfo != NULL ? 1.0 : 0.0);
      need = room > per ? (unsigned int) (nsort * per / room) + 1U : 1024U;
      gt_logger_log(logger, "B200 build: -memlimit " GT_WU " bytes -> at least %u part(s)", maximumspace, need);
    }
    if (gtb200_multi_free_memory(sfi->dev, &freeb, &totalb) == 0 && freeb > 0)
    {
      const double room = 0.92 * (double) freeb - 2.0 * tablebytes;      /* cursors mirror the leftborder table */
      const double share = nsort / (double) ndev;
      const unsigned int hbm = room > 40.0 ? (unsigned int) (share * 40.0 / room) + 1U : 1024U;
      if (hbm > 1U)
      {
        gt_logger_log(logger, "B200 build: %.1f GB of HBM free per GPU -> at least %u part(s)", (double) freeb / 1e9, hbm);
      }
      if (hbm > need)
      {
        need = hbm;
      }
    }
    if (need > numofparts)
    {
      numofparts = need;
    }
  }
  sfi->mincodes = gt_malloc(sizeof (uint64_t) * numofparts);
  sfi->maxcodes = gt_malloc(sizeof (uint64_t) * numofparts);
  sfi->widths = gt_malloc(sizeof (uint64_t) * numofparts);
  sfi->numofparts = gtb200_suftabparts(numofparts, sfi->leftborder, sfi->ci.numofallcodes, prefixlength,
                                       sfi->mincodes, sfi->maxcodes, sfi->widths);
  for (p = 0; p < sfi->numofparts; p++)
  {
    gt_logger_log(logger, "widthofpart[%d]=" GT_WU, p, (GtUword) sfi->widths[p]);
    if (largest < (GtUword) sfi->widths[p])
    {
      largest = (GtUword) sfi->widths[p];
    }
  }
  sfi->bufentries = largest;
  if (sfi->bufentries < 1024 && sfi->bufentries < sfi->ci.specialcharacters + 1)
  {
    sfi->bufentries = sfi->ci.specialcharacters + 1 < 1024 ? sfi->ci.specialcharacters + 1 : 1024;
  }
  if (sfi->bufentries == 0)
  {
    sfi->bufentries = 1;
  }
  sfi->suffixsortspace = gt_suffixsortspace_new(sfi->bufentries, sfi->totallength, sfi->suftabuint, logger);
  {
    /* Page-lock what the device copies into directly (GTB200_PIN=0 leaves everything pageable): the lcp buffer, and
       the suftab only when its entries are copied as they are (-suftabuint, several GPUs, GTB200_NARROW=0).  A GtUword suftab
       is otherwise filled by the library's host threads from a page-locked ring (narrow transfer) and can stay
       pageable - registering 8 bytes per suffix would cost more than it saves. */
    const char *pin = getenv("GTB200_PIN"), *narrow = getenv("GTB200_NARROW");
    sfi->pin = (pin == NULL || atoi(pin) != 0) && sfi->bufentries >= (GtUword) (1UL << 20);
    if (sfi->pin && (sfi->suftabuint || ndev > 1 || (narrow != NULL && atoi(narrow) == 0)))
    {
      void *raw = b200_rawptr(sfi);
      if (gtb200_host_register(raw, (size_t) sfi->bufentries * (sfi->suftabuint ? 4 : 8)) == 0)
      {
        sfi->pinned_suf = raw;
      }
    }
  }
  if (sfi->outlcpinfo != NULL)
  {
    gt_Outlcpinfo_numsuffixes2output_set(sfi->outlcpinfo, sfi->totallength + 1);
    sfi->lcpbuf = gt_malloc(sizeof (uint8_t) * (size_t) (largest + 1));
    if (sfi->pin && gtb200_host_register(sfi->lcpbuf, (size_t) (largest + 1)) == 0)
    {
      sfi->pinned_lcp = sfi->lcpbuf;
    }
    sfi->llvcap = largest + 1;
    sfi->llvbuf = gt_malloc(sizeof (gtb200_largelcp) * (size_t) sfi->llvcap);
  }
  return sfi;
}

Sfxiterator *gt_Sfxiterator_new(const GtEncseq *encseq,
                                GtReadmode readmode,
                                unsigned int prefixlength,
                                unsigned int numofparts,
                                GtUword maximumspace,
                                const Sfxstrategy *sfxstrategy,
                                GtTimer *sfxprogress,
                                bool withprogressbar,
                                GtLogger *logger,
                                GtError *err)
{
  return gt_Sfxiterator_new_withadditionalvalues(encseq, readmode, prefixlength, numofparts, maximumspace,
                                                 NULL, NULL, sfxstrategy, sfxprogress, withprogressbar,
                                                 logger, err);
}

const GtSuffixsortspace *gt_Sfxiterator_next(GtUword *numberofsuffixes, bool *specialsuffixes, Sfxiterator *sfi)
{
  if (sfi->part < sfi->numofparts)
  {
    gtb200_partinfo pi;
    const int p = sfi->part++;

    if (sfi->sfxprogress != NULL)
    {
      /* the reference's two phases of a part (src/match/sfx-suffixer.c:1964,2041) are one device pass here */
      gt_timer_show_progress(sfi->sfxprogress, "inserting suffixes into buckets", stdout);
      gt_timer_show_progress(sfi->sfxprogress, "sorting the buckets", stdout);
    }
    if (gtb200_multi_sort_part(sfi->dev, sfi->mincodes[p], sfi->maxcodes[p], b200_rawptr(sfi), sfi->suftabuint ? 1 : 0,
                         sfi->lcpbuf, sfi->llvbuf, (uint64_t) sfi->llvcap, &pi) != 0)
    {
      fprintf(stderr, "gt suffixerator (B200): %s\n", gtb200_last_error());
      exit(EXIT_FAILURE);         /* gt_Sfxiterator_next has no error channel (sfx-suffixer.h:60-62) */
    }
    if (pi.longest != UINT64_MAX)
    {
      sfi->longest = (GtUword) pi.longest;
      sfi->longest_defined = true;
    }
    if (sfi->outlcpinfo != NULL)
    {
      gt_Outlcpinfo_b200_writepart(sfi->outlcpinfo, sfi->lcpbuf, (GtUword) pi.width, sfi->llvbuf,
                                   (GtUword) pi.numoflargelcpvalues, pi.lcptabsum, (GtUword) pi.maxbranchdepth);
      if (gt_Outlcpinfo_b200_wantsdist(sfi->outlcpinfo))
      {
        /* -lcpdist: replay the regular entries of every bucket of the part, as outlcpvalues sees them
           (src/match/sfx-lcpvalues.c:622-693); the special-padded right end of a bucket is skipped */
        const uint64_t sigma = (uint64_t) gt_alphabet_num_of_chars(gt_encseq_alphabet(sfi->encseq));
        GtUword llvidx = 0;
        uint64_t code;

        for (code = sfi->mincodes[p]; code <= sfi->maxcodes[p]; code++)
        {
          const GtUword left = (GtUword) sfi->leftborder[code], right = (GtUword) sfi->leftborder[code + 1];
          GtUword padded = 0;

          if (code % sigma == sigma - 1)
          {
            padded = (GtUword) sfi->countspecialcodes[(code - (sigma - 1)) / sigma];
          }
          if (right - padded > left)
          {
            gt_Outlcpinfo_b200_distadd(sfi->outlcpinfo, sfi->lcpbuf, left - (GtUword) pi.partoffset,
                                       right - padded - (GtUword) pi.partoffset, sfi->llvbuf,
                                       (GtUword) pi.numoflargelcpvalues, (GtUword) pi.partoffset, &llvidx);
          }
        }
      }
    }
    *numberofsuffixes = (GtUword) pi.width;
    if (specialsuffixes != NULL)
    {
      *specialsuffixes = false;
    }
    return sfi->suffixsortspace;
  }
  if (sfi->tail_done < sfi->ci.specialcharacters + 1)
  {
    /* replaces gt_SSSPbuf_fillspecialnextpage: positions of the specials in ascending order, then n */
    GtUword page = sfi->ci.specialcharacters + 1 - sfi->tail_done;
    if (page > sfi->bufentries)
    {
      page = sfi->bufentries;
    }
    if (gtb200_multi_tail(sfi->dev, sfi->tail_done, page, b200_rawptr(sfi), sfi->suftabuint ? 1 : 0) != 0)
    {
      fprintf(stderr, "gt suffixerator (B200): %s\n", gtb200_last_error());
      exit(EXIT_FAILURE);
    }
    if (!sfi->longest_defined && sfi->tail_done == 0 && sfi->totallength > 0)
    {
      /* position 0 holds a special: it is the first entry of the tail */
      sfi->longest = (GtUword) sfi->ci.numofsuffixestosort;
      sfi->longest_defined = true;
    }
    sfi->tail_done += page;
    *numberofsuffixes = page;
    if (specialsuffixes != NULL)
    {
      *specialsuffixes = true;
    }
    return sfi->suffixsortspace;
  }
  return NULL;
}

int gt_Sfxiterator_bcktab2file(FILE *fp, Sfxiterator *sfi, GtError *err)
{
  gt_error_check(err);
  if (sfi->numofparts <= 1)
  {
    b200_bcktab_to_file(sfi, fp);
    gt_fa_fclose(fp);
  }
  return 0;
}

GtUword gt_Sfxiterator_longest(const Sfxiterator *sfi)
{
  return sfi->longest_defined ? sfi->longest : 0;
}
