// vtn_classify.cuh -- classify_variant's per-ALT rule (variant_manip.cpp:254-491), shared by the normalize kernels
// (vtn_normalize.cu) and the annotate_indels front end (vtn_annot.cu).
#pragma once
#include "vtn_internal.h"

// classify one ALT against REF: variant_manip.cpp:254-491.  Returns VT_* bits, clears homog when lengths differ.
static __device__ uint32_t classify_alt(const uint8_t* ref0, uint32_t rlen, const uint8_t* a, uint32_t l, bool& homog)
{
    bool lt = false, br = false;
    for (uint32_t i = 0; i < l; ++i)
    {
        uint8_t c = a[i];
        lt |= (c == '<');
        br |= (c == '[') | (c == ']');
    }
    if (lt)
    {
        uint32_t type = VTN_VT_REF;
        if (l >= 5)
        {
            if (a[0] == '<' && a[1] == 'V' && a[2] == 'N' && a[l - 1] == '>')
            {
                for (uint32_t j = 3; j < l - 1; ++j)
                    if (a[j] < '0' || a[j] > '9') type = VTN_VT_VNTR;
            }
            else if (l == 5 && a[0] == '<' && a[1] == 'S' && a[2] == 'T' && a[3] == 'R' && a[4] == '>')
                type = VTN_VT_VNTR;
            else if (a[0] == '<' && a[1] == 'S' && a[2] == 'T' && a[l - 1] == '>')
            {
                type = VTN_VT_VNTR;
                for (uint32_t j = 3; j < l - 1; ++j)
                    if ((a[j] < '0' || a[j] > '9') && a[j] != '.') type = VTN_VT_SV;
            }
        }
        return type == VTN_VT_VNTR ? VTN_VT_VNTR : VTN_VT_SV;
    }
    if (br) return VTN_VT_SV;
    if (a[0] == '.') return VTN_VT_REF;
    if (l == rlen)
    {
        bool same = true;
        for (uint32_t i = 0; i < l; ++i) same &= (a[i] == ref0[i]);
        if (same) return VTN_VT_REF;
    }
    else
        homog = false;
    // case-sensitive trim to a core (right first, then left), :360-390
    uint32_t rl = rlen, al = l, s = 0;
    while (rl != 1 && al != 1 && ref0[rl - 1] == a[al - 1]) { --rl; --al; }
    while (rl != 1 && al != 1 && ref0[s] == a[s]) { ++s; --rl; --al; }
    uint32_t mlen = rl < al ? rl : al;
    uint32_t diff = 0;
    if (mlen == 1 && rl != al)
    {
        uint8_t ls, le, ss;
        if (rl > al) { ls = ref0[s]; le = ref0[s + rl - 1]; ss = a[s]; }
        else { ls = a[s]; le = a[s + al - 1]; ss = ref0[s]; }
        if (ls != ss && le != ss) diff = 1;
    }
    else
    {
        for (uint32_t j = 0; j < mlen; ++j) diff += (ref0[s + j] != a[s + j]);
    }
    uint32_t type = 0;
    if (mlen == diff) type |= (mlen == 1) ? VTN_VT_SNP : VTN_VT_MNP;
    if (rl != al) type |= VTN_VT_INDEL;
    if (diff && diff < mlen) type |= VTN_VT_CLUMPED;
    return type;
}

