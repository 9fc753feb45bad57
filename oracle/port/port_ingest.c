/* port_ingest.c - CPU restatement of the front ends' byte-unpack routines.
 * TEST INFRASTRUCTURE (see port.h).  Pinned against the reference's own driver files
 * compiled into oracle/_ref (tests/test_oracle_port.py) and against
 * tests/golden/ingest.npz.                                                        */
#include "port.h"

/* rtlsdr_exp(), src/rcv/rtlsdr/rtlsdr.c:136-142: unsigned byte to signed, truncating */
void port_rtlsdr_exp(const unsigned char *buf, int n, signed char *out)
{
    int i;
    for (i = 0; i < n; i++) out[i] = (signed char)((buf[i] - 127.5));
}

/* stereo_exp(), src/rcv/stereo/stereo.c:163-204: dtype 1 = front end 1 (2 bit, bits 7:6),
 * dtype 2 = front end 2 (3 bit I in bits 5:3, 3 bit Q in bits 2:0)                 */
void port_stereo_exp(const unsigned char *buf, int n, int dtype, signed char *out)
{
    static const signed char base1[4] = {-3, -1, +1, +3};
    static const signed char base2[8] = {+1, +3, +5, +7, -7, -5, -3, -1};
    int i;
    if (dtype == 1) {
        for (i = 0; i < n; i++) out[i] = base1[(buf[i] >> 6) & 0x03];
    } else {
        for (i = 0; i < n; i++) {
            out[2 * i]     = base2[(buf[i] >> 3) & 0x07];
            out[2 * i + 1] = base2[buf[i] & 0x07];
        }
    }
}

/* gn3s_exp_v3(), src/rcv/gn3s/gn3s.cpp:143-177: modes x32/x38 are 2 bit real, modes
 * x36/x42 are 4 bit I/Q with I in bits {0,2} and Q in bits {1,3}; returns samples out */
int port_gn3s_exp_v3(const unsigned char *buf, int n, int i_mode, signed char *out)
{
    static const signed char lut2[4] = {1, -1, 3, -3};
    static const signed char luti[16] = {1, -1, 0, 0, 3, -3, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    static const signed char lutq[16] = {1, 0, -1, 0, 0, 0, 0, 0, 3, 0, -3, 0, 0, 0, 0, 0};
    int i, k = 0, t = 0;
    if ((i_mode % 100) == 32 || (i_mode % 100) == 38) t = 2;
    if ((i_mode % 100) == 36 || (i_mode % 100) == 42) t = 4;
    for (i = 0; i < n; i++) {
        if (t == 2) out[k++] = lut2[buf[i] & 0x03];
        else if (t == 4) { out[k++] = luti[buf[i] & 0x05]; out[k++] = lutq[buf[i] & 0x0A]; }
    }
    return k;
}
