/*
 * oracle/shims/fitsio.h -- TEST INFRASTRUCTURE, not product code.
 *
 * Stand-in for the nine CFITSIO entry points the reference uses
 * (src/FITSWriter.cpp:44-148: create_file, close_file, create_img,
 * write_key, write_comment, write_date, write_pix; src/FITSWriter.h:31,39:
 * get_errstatus, read_errmsg).  CFITSIO is not installed in this image.
 * The shim writes standard, UNCOMPRESSED primary-HDU FITS files (a
 * "[compress]" suffix on the file name is accepted and ignored: tile
 * compression is unavailable on this box).
 */
#ifndef RO_ORACLE_FITSIO_SHIM_H
#define RO_ORACLE_FITSIO_SHIM_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ro_shim_fitsfile fitsfile;

#define FLEN_FILENAME 1025
#define FLEN_KEYWORD  75
#define FLEN_CARD     81
#define FLEN_VALUE    71
#define FLEN_COMMENT  73
#define FLEN_ERRMSG   81
#define FLEN_STATUS   31

#define BYTE_IMG      8
#define SHORT_IMG    16
#define LONG_IMG     32
#define LONGLONG_IMG 64
#define FLOAT_IMG   -32
#define DOUBLE_IMG  -64

#define TSTRING     16
#define TSHORT      21
#define TINT        31
#define TLONG       41
#define TFLOAT      42
#define TLONGLONG   81
#define TDOUBLE     82

#define FILE_NOT_CREATED 105
#define WRITE_ERROR      106
#define BAD_DATATYPE     410
#define NULL_INPUT_PTR   115

int fits_create_file(fitsfile **fptr, const char *filename, int *status);
int fits_close_file(fitsfile *fptr, int *status);
int fits_create_img(fitsfile *fptr, int bitpix, int naxis, long *naxes, int *status);
int fits_write_key(fitsfile *fptr, int datatype, const char *keyname,
                   void *value, const char *comment, int *status);
int fits_write_comment(fitsfile *fptr, const char *comment, int *status);
int fits_write_date(fitsfile *fptr, int *status);
int fits_write_pix(fitsfile *fptr, int datatype, long *firstpix,
                   long long nelem, void *array, int *status);
void fits_get_errstatus(int status, char *errtext);
int fits_read_errmsg(char *errmsg);

#ifdef __cplusplus
}
#endif

#endif
