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
