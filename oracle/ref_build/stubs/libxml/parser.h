// TEST INFRASTRUCTURE (oracle/): libxml2 is not in this image.  The reference's Core/XmlParser.hh includes
// <libxml/parser.h> and mentions the type names below in member declarations; Am, Speech and Search headers include it
// in turn.  Declaring the names (opaque, no functions) lets those HEADERS parse.  No translation unit that calls
// libxml2 is compiled (oracle/ref_build/Makefile leaves Core/XmlParser.cc and Fsa's XML readers out).
#pragma once
typedef unsigned char xmlChar;
struct _xmlParserInput;
struct _xmlEntity;
struct _xmlParserCtxt;
struct _xmlEnumeration;
struct _xmlElementContent;
struct _xmlSAXHandler {
    int unused;
};
typedef _xmlParserInput*    xmlParserInputPtr;
typedef _xmlEntity*         xmlEntityPtr;
typedef _xmlParserCtxt*     xmlParserCtxtPtr;
typedef _xmlEnumeration*    xmlEnumerationPtr;
typedef _xmlElementContent* xmlElementContentPtr;
typedef _xmlSAXHandler      xmlSAXHandler;
typedef xmlSAXHandler*      xmlSAXHandlerPtr;
