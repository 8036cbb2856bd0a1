"""``distance((lat, lon), (lat, lon)).meters`` on the WGS-84 ellipsoid (Vincenty's inverse formula, which
agrees with geopy's default geodesic to well below a metre; the reference rounds the resulting link
delay to whole milliseconds, coordsim/reader/reader.py:223-225)."""
import math

_A, _F = 6378137.0, 1 / 298.257223563
_B = (1 - _F) * _A


class distance:
    def __init__(self, p1, p2):
        self.meters = _vincenty(float(p1[0]), float(p1[1]), float(p2[0]), float(p2[1]))

    @property
    def km(self):
        return self.meters / 1000.0

    kilometers = km

    @property
    def m(self):
        return self.meters


geodesic = distance


def _vincenty(lat1, lon1, lat2, lon2):
    if lat1 == lat2 and lon1 == lon2:
        return 0.0
    U1 = math.atan((1 - _F) * math.tan(math.radians(lat1)))
    U2 = math.atan((1 - _F) * math.tan(math.radians(lat2)))
    L = math.radians(lon2 - lon1)
    sU1, cU1, sU2, cU2 = math.sin(U1), math.cos(U1), math.sin(U2), math.cos(U2)
    lam = L
    for _ in range(200):
        sl, cl = math.sin(lam), math.cos(lam)
        ss = math.sqrt((cU2 * sl) ** 2 + (cU1 * sU2 - sU1 * cU2 * cl) ** 2)
        if ss == 0:
            return 0.0
        cs = sU1 * sU2 + cU1 * cU2 * cl
        sig = math.atan2(ss, cs)
        sa = cU1 * cU2 * sl / ss
        c2a = 1 - sa * sa
        c2sm = cs - 2 * sU1 * sU2 / c2a if c2a != 0 else 0.0
        C = _F / 16 * c2a * (4 + _F * (4 - 3 * c2a))
        new = L + (1 - C) * _F * sa * (sig + C * ss * (c2sm + C * cs * (-1 + 2 * c2sm * c2sm)))
        if abs(new - lam) < 1e-12:
            lam = new
            break
        lam = new
    u2 = c2a * (_A * _A - _B * _B) / (_B * _B)
    Ak = 1 + u2 / 16384 * (4096 + u2 * (-768 + u2 * (320 - 175 * u2)))
    Bk = u2 / 1024 * (256 + u2 * (-128 + u2 * (74 - 47 * u2)))
    ds = Bk * ss * (c2sm + Bk / 4 * (cs * (-1 + 2 * c2sm * c2sm) - Bk / 6 * c2sm * (-3 + 4 * ss * ss) * (-3 + 4 * c2sm * c2sm)))
    return _B * Ak * (sig - ds)
