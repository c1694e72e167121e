/*
 * integration/sfx_lcpvalues_b200.c -- B200 build of src/match/sfx-lcpvalues.c.
 *
 * GtOutlcpinfo and its file handles are private to that translation unit, so the B200 build
 * compiles the UNMODIFIED reference file (included below from where it lies, nothing copied) and
 * adds ONE function: the writer the device-backed Sfxiterator calls once per part.  The device
 * already delivers final .lcp bytes (255 = overflow) and .llv records in ascending index order, so
 * the function appends them exactly where outlcpvalues (src/match/sfx-lcpvalues.c:371-433) and
 * outsmalllcpvalues (:113-123) would, and keeps the counters the .prj writer reads
 * (gt_Outlcpinfo_numoflargelcpvalues / _lcptabsum / _maxbranchdepth) up to date.
 * gt_Outlcpinfo_delete (:453-503) then pads the tail with zeros as before.
 * -lcpdist: the histogram outlcpvalues fills value by value (:415-418) covers the entries that pass
 * through it, i.e. the regular suffixes of every bucket; gt_Outlcpinfo_b200_distadd replays those
 * entries from the part's .lcp bytes and .llv records (the caller walks the bucket table).
 */
#include "match/sfx-lcpvalues.c"

void gt_Outlcpinfo_b200_writepart(GtOutlcpinfo *outlcpinfo, const uint8_t *lcp, GtUword width,
                                  const void *llv, GtUword nllv, double lcptabsum, GtUword maxbranchdepth)
{
  Lcpoutput2file *l2f;

  gt_assert(outlcpinfo != NULL);
  l2f = outlcpinfo->lcpsubtab.lcp2file;
  gt_assert(l2f != NULL);
  if (width > 0)
  {
    gt_xfwrite(lcp, sizeof (uint8_t), (size_t) width, l2f->outfplcptab);
  }
  if (nllv > 0)
  {
    gt_xfwrite(llv, sizeof (Largelcpvalue), (size_t) nllv, l2f->outfpllvtab);
  }
  l2f->countoutputlcpvalues += width;
  l2f->totalnumoflargelcpvalues += nllv;
  outlcpinfo->lcpsubtab.lcptabsum += lcptabsum;
  if (l2f->maxbranchdepth < maxbranchdepth)
  {
    l2f->maxbranchdepth = maxbranchdepth;
  }
}

bool gt_Outlcpinfo_b200_wantsdist(const GtOutlcpinfo *outlcpinfo)
{
  return outlcpinfo != NULL && outlcpinfo->lcpsubtab.distlcpvalues != NULL;
}

/* entries [from,to) of the part (part-local indices) went through outlcpvalues in the reference;
   *llvidx walks the part's .llv records, which are in ascending index order */
void gt_Outlcpinfo_b200_distadd(GtOutlcpinfo *outlcpinfo, const uint8_t *lcp, GtUword from, GtUword to,
                                const void *llv, GtUword nllv, GtUword partoffset, GtUword *llvidx)
{
  const Largelcpvalue *large = (const Largelcpvalue *) llv;
  GtUword idx;

  gt_assert(outlcpinfo != NULL && outlcpinfo->lcpsubtab.distlcpvalues != NULL);
  for (idx = from; idx < to; idx++)
  {
    if (lcp[idx] < (uint8_t) LCPOVERFLOW)
    {
      gt_disc_distri_add(outlcpinfo->lcpsubtab.distlcpvalues, (GtUword) lcp[idx]);
    } else
    {
      while (*llvidx < nllv && large[*llvidx].position < partoffset + idx)
      {
        (*llvidx)++;
      }
      gt_assert(*llvidx < nllv && large[*llvidx].position == partoffset + idx);
      gt_disc_distri_add(outlcpinfo->lcpsubtab.distlcpvalues, large[*llvidx].value);
    }
  }
}
