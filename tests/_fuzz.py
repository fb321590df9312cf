"""Adversarial bodies for the scan stage: fragments of every needle in both cases, code points whose
upper() contains ASCII, digits of other scripts, NBSP / bullet, card and date forms packed together."""
import random

PIECES = ["OTP", "otp", "Otp", "CODE:", "code:", "PASS:", "PASS=", "pass=", "NOT ENOUGH FUNDS", "not enough funds",
          "INSUFFICIENT FUNDS", "insuﬃcient funds", "CREDIT PAYMENT", "credit payment", "C2C RECEIVED",
          "PERSON TO PERSON", "person to perſon", "Daily limit exceeded", "Daily limit exceeded:",
          "daily limit exceeded", "DAILY LIMIT EXCEEDED", "Daily  limit exceeded", "PAſſ:", "paß=",
          "ınsufficient", "CREDİT", "ŉot enough", "ﬁ", "ﬂ", "ﬅ", "ǰ", "ẖ",
          "4083***7538", "4083•••7538", "1234***5678***9012", "12345***67890", "1234****5678",
          "٤٠٨٣***٧٥٣٨", "40٨٣***75٣٨", "***", "*", "•",
          " ", " ", "\n", ".", ":", "=", "06.05.25", "10.06.2025", "31.02.2024", "29.02.24", "00.01.2024",
          "1٢.0١.2025", "١٢.01.2025", "12.01.٢٠٢٥", "99.99.99", "1.1.2025",
          "01.13.2025", "32.01.25", "0", "7", "42", "2025", "A", "z", "я", "€", "\U0001F600", "O", "T", "P",
          "C", "OD", "E:", "PAS", "S:", "Dai", "ly limit", " exceeded",
          # scripts whose lead bytes are not special (they stay in the scan's word loop), next to needles and forms
          "Покупка", "Баланс", "отп", "ОТП", "Գնում", "ΑΓΟΡΑ", "中文", "日本", "é", "Ü", "ö", "я1", "1я",
          "я06.05.25", "06.05.25я", "я4083***7538я", "OTя", "OяTP", "Ωcode:", "ЖPASS=Ж"]


def bodies(n, seed):
    rng = random.Random(seed)
    out = []
    for _ in range(n):
        k = rng.choice([1, 2, 3, 5, 8, 13, 30])
        s = "".join(rng.choice(PIECES) for _ in range(k))
        if rng.random() < 0.3:                      # cut somewhere: partial needles at the ends
            a = rng.randrange(len(s) + 1)
            s = s[a:] if rng.random() < 0.5 else s[:a]
        out.append(s)
    return out
