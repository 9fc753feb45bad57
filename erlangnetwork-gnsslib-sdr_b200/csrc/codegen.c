/* codegen.c - ranging-code tables for the code-spectrum bank (host C).
 *
 * Replaces, for the hot path's needs, the reference's gencode() dispatch
 * (src/sdrcode.c:523-539) for CTYPE_L1CA / CTYPE_L1SBAS (src/sdrcode.c:101-154)
 * and adds the GLONASS G1 case that the reference defines (src/sdrcode.c:426-444)
 * but never dispatches to.
 *
 * C/A code: G1 = x^10+x^3+1, G2 = x^10+x^9+x^8+x^6+x^3+x^2+1, both seeded with
 * all ones; chip i of PRN p is G1[i] xor G2[(i - delay[p]) mod 1023]
 * (IS-GPS-200, RTCA DO-229 and IS-QZSS delay assignments).  Output uses the
 * reference's polarity: logic 0 -> -1 ... the reference emits
 * code = -G1*G2 in +-1 arithmetic with registers initialised to -1, i.e.
 * (g1 xor g2) == 1  ->  +1 ... see c/a_chip() below, verified chip-for-chip
 * against the reference generator: tests/test_oracle_port.py (all 210 PRNs and G1 against
 * the reference build), tests/golden/codes.npz (the same on machines without the reference) and
 * tests/test_gpu_primitives.py::test_gencode_on_device (the device twin against this file).
 */
#include <stdint.h>
#include <string.h>
#include "sdrb200_internal.h"

/* G2 code delay in chips for PRN 1..210 */
static const uint16_t k_g2_delay[210] = {
    5,6,7,8,17,18,139,140,141,251, 252,254,255,256,257,258,469,470,471,472,
    473,474,509,512,513,514,515,516,859,860, 861,862,863,950,947,948,950,67,103,91,
    19,679,225,625,946,638,161,1001,554,280, 710,709,775,864,558,220,397,55,898,759,
    367,299,1018,729,695,780,801,788,732,34, 320,327,389,407,525,405,221,761,260,326,
    955,653,699,422,188,438,959,539,879,677, 586,153,792,814,446,264,1015,278,536,819,
    156,957,159,712,885,461,248,713,126,807, 279,122,197,693,632,771,467,647,203,145,
    175,52,21,237,235,886,657,634,762,355, 1012,176,603,130,359,595,68,386,797,456,
    499,883,307,127,211,121,118,163,628,853, 484,289,811,202,1021,463,568,904,670,230,
    911,684,309,644,932,12,314,891,212,185, 675,503,150,395,345,846,798,992,357,995,
    877,112,144,476,193,109,445,291,87,399, 292,901,339,208,711,189,263,537,663,942,
    173,900,30,500,935,556,373,85,652,310
};

const uint16_t *sdrb200_g2_delay_table(void) { return k_g2_delay; }

/* run a 10-stage Fibonacci LFSR; taps is a bit mask over stages 1..10 (bit s-1) */
static void lfsr10(uint32_t taps, uint8_t *out /*1023*/)
{
    uint32_t reg = 0x3FF;                       /* all ones */
    int i;
    for (i = 0; i < 1023; i++) {
        uint32_t fb = (uint32_t)__builtin_parity(reg & taps);
        out[i] = (uint8_t)((reg >> 9) & 1);     /* stage 10 is the output */
        reg = ((reg << 1) | fb) & 0x3FF;
    }
}

int sdrb200_gencode_l1ca(int prn, int8_t *code /*1023*/)
{
    uint8_t g1[1023], g2[1023];
    int i, d;
    if (prn < 1 || prn > 210 || !code) return -1;
    lfsr10((1u << 2) | (1u << 9), g1);                                   /* stages 3,10        */
    lfsr10((1u << 1) | (1u << 2) | (1u << 5) | (1u << 7) | (1u << 8) | (1u << 9), g2); /* 2,3,6,8,9,10 */
    d = k_g2_delay[prn - 1];
    for (i = 0; i < 1023; i++) {
        int chip = g1[i] ^ g2[(i + 1023 - d) % 1023];
        /* reference polarity: registers hold -1 for logic 1, code = -G1*G2 */
        code[i] = chip ? 1 : -1;
    }
    return 1023;
}

/* GLONASS ST code: x^9+x^5+1, output from stage 7, 511 chips */
int sdrb200_gencode_g1(int8_t *code /*511*/)
{
    uint32_t reg = 0x1FF;
    int i;
    if (!code) return -1;
    for (i = 0; i < 511; i++) {
        uint32_t fb = ((reg >> 4) ^ (reg >> 8)) & 1;
        code[i] = ((reg >> 6) & 1) ? 1 : -1;
        reg = ((reg << 1) | fb) & 0x1FF;
    }
    return 511;
}

int sdrb200_gencode(int prn, int ctype, int8_t *code, double *crate)
{
    switch (ctype) {
    case SDRB200_CTYPE_L1CA:
    case SDRB200_CTYPE_L1SBAS:
        if (crate) *crate = 1.023e6;
        return sdrb200_gencode_l1ca(prn, code);
    case SDRB200_CTYPE_G1:
        if (crate) *crate = 0.511e6;
        return sdrb200_gencode_g1(code);
    default:
        return -1;
    }
}
