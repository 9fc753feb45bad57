/* Empty stand-in for <libusb-1.0/libusb.h>.
 * The reference's sdr.h includes this header unconditionally (sdr.h:79) but the
 * hot path never touches USB.  This stub exists only so that the oracle build
 * can compile the reference sources where they lie; it is test infrastructure. */
#ifndef ORACLE_STUB_LIBUSB_H
#define ORACLE_STUB_LIBUSB_H
#endif
