/*
 * oracle/shims/fitsio_shim.cpp -- TEST INFRASTRUCTURE, not product code.
 *
 * Minimal writer behind oracle/shims/fitsio.h.  Header cards follow the
 * FITS standard fixed format and CFITSIO's value formatting conventions
 * (TFLOAT -> "%.7G", TDOUBLE -> "%.15G", a '.' appended when the text has
 * neither '.' nor 'E'; strings quoted and blank-padded to 8 characters;
 * numeric values right-justified to column 30).  Pixel data are stored
 * big-endian after the header, padded to 2880-byte blocks.
 */
#include "fitsio.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>
#include <stdint.h>

struct ro_shim_fitsfile {
	std::string path;
	bool compress_requested;
	std::vector<std::string> cards;
	int bitpix;
	long naxes[2];
	std::vector<unsigned char> data; /* big-endian pixel bytes */
};

static std::string pad80(const std::string &s)
{
	std::string c = s.substr(0, 80);
	c.resize(80, ' ');
	return c;
}

static std::string card_value(const char *key, const std::string &value, const char *comment)
{
	char buf[128];
	std::string v = value;
	if (v.size() < 20) v = std::string(20 - v.size(), ' ') + v; /* right-justify to col 30 */
	snprintf(buf, sizeof buf, "%-8.8s= %s", key, v.c_str());
	std::string c = buf;
	if (comment && comment[0]) { c += " / "; c += comment; }
	return pad80(c);
}

static std::string card_string(const char *key, const char *value, const char *comment)
{
	std::string q = "'";
	for (const char *p = value; *p; ++p) { q += *p; if (*p == '\'') q += '\''; }
	while (q.size() < 9) q += ' ';
	q += "'";
	char buf[160];
	snprintf(buf, sizeof buf, "%-8.8s= %-20s", key, q.c_str());
	std::string c = buf;
	if (comment && comment[0]) { c += " / "; c += comment; }
	return pad80(c);
}

static std::string fmt_real(double v, int digits)
{
	char buf[64];
	snprintf(buf, sizeof buf, "%.*G", digits, v);
	if (!strchr(buf, '.') && strchr(buf, 'E')) snprintf(buf, sizeof buf, "%.1E", v);
	if (!strchr(buf, '.') && !strchr(buf, 'E')) strcat(buf, ".");
	return buf;
}

extern "C" int fits_create_file(fitsfile **fptr, const char *filename, int *status)
{
	if (*status > 0) return *status;
	std::string name = filename ? filename : "";
	bool clobber = false;
	if (!name.empty() && name[0] == '!') { clobber = true; name.erase(0, 1); }
	bool compress = false;
	size_t br = name.find('[');
	if (br != std::string::npos) {
		compress = name.compare(br, 9, "[compress") == 0;
		name.erase(br);
	}
	if (!clobber) {
		FILE *t = fopen(name.c_str(), "rb");
		if (t) { fclose(t); return *status = FILE_NOT_CREATED; }
	}
	FILE *f = fopen(name.c_str(), "wb");
	if (!f) return *status = FILE_NOT_CREATED;
	fclose(f);
	ro_shim_fitsfile *ff = new ro_shim_fitsfile;
	ff->path = name;
	ff->compress_requested = compress;
	ff->bitpix = 0;
	ff->naxes[0] = ff->naxes[1] = 0;
	*fptr = ff;
	return *status;
}

extern "C" int fits_create_img(fitsfile *f, int bitpix, int naxis, long *naxes, int *status)
{
	if (*status > 0) return *status;
	if (!f || naxis != 2) return *status = NULL_INPUT_PTR;
	f->bitpix = bitpix;
	f->naxes[0] = naxes[0];
	f->naxes[1] = naxes[1];
	char v[32];
	f->cards.push_back(card_value("SIMPLE", "T", "file does conform to FITS standard"));
	snprintf(v, sizeof v, "%d", bitpix);
	f->cards.push_back(card_value("BITPIX", v, "number of bits per data pixel"));
	f->cards.push_back(card_value("NAXIS", "2", "number of data axes"));
	snprintf(v, sizeof v, "%ld", naxes[0]);
	f->cards.push_back(card_value("NAXIS1", v, "length of data axis 1"));
	snprintf(v, sizeof v, "%ld", naxes[1]);
	f->cards.push_back(card_value("NAXIS2", v, "length of data axis 2"));
	f->cards.push_back(card_value("EXTEND", "T", "FITS dataset may contain extensions"));
	f->cards.push_back(pad80("COMMENT   FITS (Flexible Image Transport System) format is defined in 'Astronomy"));
	f->cards.push_back(pad80("COMMENT   and Astrophysics', volume 376, page 359; bibcode: 2001A&A...376..359H"));
	size_t bytes = (size_t)(abs(bitpix) / 8) * (size_t)naxes[0] * (size_t)naxes[1];
	f->data.assign(bytes, 0);
	return *status;
}

extern "C" int fits_write_key(fitsfile *f, int datatype, const char *keyname,
                              void *value, const char *comment, int *status)
{
	if (*status > 0) return *status;
	if (!f) return *status = NULL_INPUT_PTR;
	char v[64];
	switch (datatype) {
	case TSTRING:
		f->cards.push_back(card_string(keyname, (const char *)value, comment));
		break;
	case TFLOAT:
		f->cards.push_back(card_value(keyname, fmt_real((double)*(float *)value, 7), comment));
		break;
	case TDOUBLE:
		f->cards.push_back(card_value(keyname, fmt_real(*(double *)value, 15), comment));
		break;
	case TINT:
		snprintf(v, sizeof v, "%d", *(int *)value);
		f->cards.push_back(card_value(keyname, v, comment));
		break;
	case TLONG:
		snprintf(v, sizeof v, "%ld", *(long *)value);
		f->cards.push_back(card_value(keyname, v, comment));
		break;
	case TLONGLONG:
		snprintf(v, sizeof v, "%lld", *(long long *)value);
		f->cards.push_back(card_value(keyname, v, comment));
		break;
	default:
		return *status = BAD_DATATYPE;
	}
	return *status;
}

extern "C" int fits_write_comment(fitsfile *f, const char *comment, int *status)
{
	if (*status > 0) return *status;
	if (!f) return *status = NULL_INPUT_PTR;
	std::string s = comment ? comment : "";
	do {
		f->cards.push_back(pad80("COMMENT " + s.substr(0, 72)));
		s.erase(0, s.size() < 72 ? s.size() : 72);
	} while (!s.empty());
	return *status;
}

extern "C" int fits_write_date(fitsfile *f, int *status)
{
	if (*status > 0) return *status;
	if (!f) return *status = NULL_INPUT_PTR;
	time_t t = time(NULL);
	struct tm g;
	gmtime_r(&t, &g);
	char buf[32];
	strftime(buf, sizeof buf, "%Y-%m-%dT%H:%M:%S", &g);
	f->cards.push_back(card_string("DATE", buf, "file creation date (YYYY-MM-DDThh:mm:ss UT)"));
	return *status;
}

extern "C" int fits_write_pix(fitsfile *f, int datatype, long *firstpix,
                              long long nelem, void *array, int *status)
{
	if (*status > 0) return *status;
	if (!f) return *status = NULL_INPUT_PTR;
	const size_t esz = (size_t)(abs(f->bitpix) / 8);
	const size_t first = (size_t)(firstpix[0] - 1) + (size_t)(firstpix[1] - 1) * (size_t)f->naxes[0];
	if ((first + (size_t)nelem) * esz > f->data.size()) return *status = WRITE_ERROR;
	unsigned char *dst = &f->data[first * esz];
	if (datatype == TFLOAT && f->bitpix == FLOAT_IMG) {
		const unsigned char *src = (const unsigned char *)array;
		for (long long i = 0; i < nelem; i++)
			for (int b = 0; b < 4; b++) dst[4 * i + b] = src[4 * i + (3 - b)];
	} else if (datatype == TSHORT && f->bitpix == SHORT_IMG) {
		const unsigned char *src = (const unsigned char *)array;
		for (long long i = 0; i < nelem; i++) { dst[2 * i] = src[2 * i + 1]; dst[2 * i + 1] = src[2 * i]; }
	} else {
		return *status = BAD_DATATYPE;
	}
	return *status;
}

extern "C" int fits_close_file(fitsfile *f, int *status)
{
	if (!f) return *status;
	FILE *out = fopen(f->path.c_str(), "wb");
	if (!out) { delete f; return *status = WRITE_ERROR; }
	std::string hdr;
	for (size_t i = 0; i < f->cards.size(); i++) hdr += f->cards[i];
	hdr += pad80("END");
	while (hdr.size() % 2880) hdr += ' ';
	fwrite(hdr.data(), 1, hdr.size(), out);
	if (!f->data.empty()) fwrite(&f->data[0], 1, f->data.size(), out);
	size_t rem = f->data.size() % 2880;
	if (rem) { std::vector<unsigned char> z(2880 - rem, 0); fwrite(&z[0], 1, z.size(), out); }
	fclose(out);
	delete f;
	return *status;
}

extern "C" void fits_get_errstatus(int status, char *errtext)
{
	snprintf(errtext, FLEN_STATUS, "shim status %d", status);
}

extern "C" int fits_read_errmsg(char *errmsg)
{
	errmsg[0] = 0;
	return 0;
}
